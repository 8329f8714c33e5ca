"""Pure-PyTorch gather/scatter for batch-first decoders (reference: data_routing/utils.py:9-27)."""
import torch


def bsh_decoder_gather(reserved_length, hidden_states, mask):
    b, s, h = hidden_states.shape
    rand = [torch.randperm(s, device=hidden_states.device)[:reserved_length].sort().values for _ in range(b)]
    idx = torch.stack(rand)
    part = torch.gather(hidden_states, 1, idx[:, :, None].expand(b, reserved_length, h))
    part_mask = mask[:, :, :reserved_length, :reserved_length] if mask is not None else None
    return part, idx, part_mask


def bsh_decoder_scatter(hidden_states, part_hidden_states, rand_list):
    b, k, h = part_hidden_states.shape
    return hidden_states.scatter(1, rand_list[:, :, None].expand(b, k, h), part_hidden_states)
