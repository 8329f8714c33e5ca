"""Cross entropy with logits sharded along the sequence (reference: deepspeed/sequence/cross_entropy.py:13-58):
each rank computes the per-token loss of its ``[S/P, B]`` slice, the slices are all-gathered to ``[S, B]``;
backward hands every rank the gradient slice of its own tokens."""
import torch

from ..comm import comm as dist


class _VocabSequenceParallelCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vocab_seq_parallel_logits, target, sp_group):
        logits = vocab_seq_parallel_logits.float()
        logits = logits - logits.max(dim=-1, keepdim=True)[0]
        exp = logits.exp()
        sum_exp = exp.sum(dim=-1, keepdim=True)
        softmax = exp / sum_exp
        picked = logits.gather(-1, target.unsqueeze(-1)).squeeze(-1)
        loss = sum_exp.squeeze(-1).log() - picked                       # [S/P, B]
        ctx.save_for_backward(softmax, target)
        ctx.seqlen = logits.size(0) * dist.get_world_size(sp_group)
        ctx.sp_world, ctx.sp_rank = dist.get_world_size(sp_group), dist.get_rank(sp_group)
        if ctx.sp_world == 1:
            return loss
        out = torch.empty((ctx.seqlen,) + tuple(loss.shape[1:]), dtype=loss.dtype, device=loss.device)
        dist.all_gather_into_tensor(out, loss.contiguous(), group=sp_group)
        return out

    @staticmethod
    def backward(ctx, grad_output):
        softmax, target = ctx.saved_tensors
        step = ctx.seqlen // ctx.sp_world
        g = grad_output[ctx.sp_rank * step: (ctx.sp_rank + 1) * step]
        grad = softmax.clone()
        grad.scatter_add_(-1, target.unsqueeze(-1), -torch.ones_like(target, dtype=grad.dtype).unsqueeze(-1))
        grad.mul_(g.unsqueeze(-1))
        return grad, None, None


def vocab_sequence_parallel_cross_entropy(vocab_parallel_logits, target, sp_group):
    return _VocabSequenceParallelCrossEntropy.apply(vocab_parallel_logits, target, sp_group)
