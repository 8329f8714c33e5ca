"""Compression layers: masks, dimension reduction, fake quantisation (reference behaviour:
``deepspeed/compression/basic_layer.py``; oracle = the dense computation written out by hand)."""
import pytest
import torch
from torch import nn

from shuffle_exchange_b200.compression.basic_layer import (BNLayer_Compress, Conv2dLayer_Compress, Embedding_Compress,
                                                           LinearLayer_Compress, QuantAct)


def test_linear_sparse_l1_and_fix():
    torch.manual_seed(0)
    lin = LinearLayer_Compress(16, 8)
    w0 = lin.weight.data.clone()
    lin.enable_sparse_pruning(0.25, "l1")
    assert "sparse_pruning_mask" in dict(lin.named_buffers())
    x = torch.randn(3, 16)
    dense = lin(x)
    lin.sparse_pruning_enabled = True
    mask = lin.get_mask("sparse")
    assert int(mask.sum()) == 32                             # 25% of 128 kept
    assert mask.flatten()[w0.abs().flatten().argmax()]
    torch.testing.assert_close(lin(x), torch.nn.functional.linear(x, w0 * mask, lin.bias))
    assert not torch.allclose(lin(x), dense)
    lin.fix_sparse_pruning_helper()
    assert lin.sparse_pruning_method is None and not hasattr(lin, "sparse_pruning_mask")
    torch.testing.assert_close(lin.weight.data, w0 * mask)


def test_linear_row_pruning_dim_reduction_feeds_next_layer():
    torch.manual_seed(1)
    a, b = LinearLayer_Compress(8, 12), LinearLayer_Compress(12, 4)
    a.enable_row_pruning(0.5, "l1")
    a.row_pruning_enabled = True
    x = torch.randn(5, 8)
    want = b(a(x))                                           # masked rows of a are zero -> columns of b are unused
    mask = a.fix_row_col_pruning_helper(dim_reduction=True)
    assert a.out_features == 6 and a.weight.shape == (6, 8) and a.bias.shape == (6,)
    assert b.fix_row_col_pruning_helper(mask, dim_reduction=True) is None
    assert b.in_features == 6
    torch.testing.assert_close(b(a(x)), want)


def test_linear_row_topk_scores_get_gradients():
    lin = LinearLayer_Compress(8, 8)
    lin.enable_row_pruning(0.5, "topk")
    lin.row_pruning_enabled = True
    assert lin.row_mask_scores.shape == (8, 1)
    lin(torch.randn(2, 8)).sum().backward()
    assert lin.row_mask_scores.grad is not None and lin.row_mask_scores.grad.abs().sum() > 0
    mask = lin.fix_row_col_pruning_helper()
    assert int(mask.sum()) == 4 and not hasattr(lin, "row_mask_scores")
    assert int((lin.weight.data.abs().sum(dim=1) == 0).sum()) == 4


def test_head_pruning_output_and_qkv_projection():
    torch.manual_seed(2)
    heads, hd, hidden = 4, 4, 16
    qkv, out = LinearLayer_Compress(hidden, hidden), LinearLayer_Compress(hidden, hidden)
    out.enable_head_pruning(0.5, "topk", heads)
    out.head_pruning_enabled = True
    with pytest.raises(NotImplementedError):
        LinearLayer_Compress(4, 4).enable_head_pruning(0.5, "l1", 2)
    x = torch.randn(3, hidden)
    want = out(qkv(x))
    mask = out.fix_head_pruning_helper(dim_reduction=True)
    assert out.in_features == 2 * hd
    qkv.fix_head_pruning_helper(mask, heads, dim_reduction=True)
    assert qkv.out_features == 2 * hd
    torch.testing.assert_close(out(qkv(x)), want)


def test_weight_quantisation_bit_widths_and_fix():
    lin = LinearLayer_Compress(32, 4)
    for bits, levels in ((8, None), (2, 3), (1, 2)):
        lin.enable_weight_quantization(bits, bits, 1, True, "symmetric", 1)
        lin.weight_quantization_enabled = True
        q = lin._forward_weight()
        if levels:
            assert q.unique().numel() <= levels
        else:
            assert (q - lin.weight).abs().max() <= lin.weight.abs().max() / 100
    with pytest.raises(AssertionError):
        lin.enable_weight_quantization(2, 2, 1, True, "asymmetric", 1)
    lin.enable_weight_quantization(4, 4, 1, True, "asymmetric", 4)
    lin.fix_weight_quantization()
    assert lin.weight_quantization_enabled_in_forward is False
    assert lin.weight.data.reshape(4, -1)[0].unique().numel() <= 16


def test_activation_quantisation_dynamic_and_static():
    lin = LinearLayer_Compress(16, 16)
    lin.enable_activation_quantization(8, "symmetric", "dynamic")
    lin.activation_quantization_enabled = True
    x = torch.randn(2, 5, 16)
    ref = torch.nn.functional.linear(x, lin.weight, lin.bias)
    assert (lin(x) - ref).abs().max() < 0.1
    lin.enable_activation_quantization(8, "asymmetric", "static")
    assert isinstance(lin.activation_quantizer, QuantAct)
    lin.train()
    lin(x)
    lo, hi = lin.activation_quantizer.x_min_max.tolist()
    assert lo < 0 < hi
    lin.eval()
    lin(10 * x)                                              # frozen in eval
    assert lin.activation_quantizer.x_min_max.tolist() == [lo, hi]
    with pytest.raises(AssertionError):
        lin.enable_activation_quantization(3, "symmetric", "dynamic")


def test_conv_channel_pruning_with_batchnorm():
    torch.manual_seed(3)
    c1, bn, c2 = Conv2dLayer_Compress(3, 8, 3), BNLayer_Compress(8), Conv2dLayer_Compress(8, 4, 1)
    bn.eval()
    c1.enable_channel_pruning(0.5, "l1")
    c1.channel_pruning_enabled = True
    x = torch.randn(2, 3, 6, 6)
    with torch.no_grad():
        bn.bias.zero_()                                      # pruned channels must contribute nothing after the norm
        bn.running_mean.zero_()
        want = c2(bn(c1(x)))
        mask = c1.fix_channel_pruning_helper(dim_reduction=True)
        bn.fix_channel_pruning_helper(mask)
        c2.fix_channel_pruning_helper(mask, dim_reduction=True)
        assert (c1.out_channels, bn.num_features, c2.in_channels) == (4, 4, 4)
        torch.testing.assert_close(c2(bn(c1(x))), want, atol=1e-5, rtol=1e-5)
    with pytest.raises(AssertionError):
        c1.enable_activation_quantization(8, "symmetric", "dynamic")


def test_embedding_quantises_per_row():
    emb = Embedding_Compress(10, 8)
    emb.enable_weight_quantization(4, 4, 1, True, "symmetric", 1)
    assert emb.weight_quantize_num_groups == 10
    emb.weight_quantization_enabled = True
    out = emb(torch.tensor([1, 2]))
    assert out[0].unique().numel() <= 16
    emb.fix_weight_quantization()
    assert emb.weight.data[3].unique().numel() <= 16
