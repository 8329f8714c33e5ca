// ragged_host.cpp — host-side schedulers of the ragged inference engine.
//
//   sxbh_build_atoms        split every in-flight sequence into q-blocks ("attention atoms") with the KV extent each
//                           block attends to (reference inference/v2/kernels/ragged_ops/atom_builder/atom_builder.cpp)
//   sxbh_fill_ragged_batch  build token->sequence / token->position / last-token tables for a batch in one pass
//
// Plain C ABI, no torch / CUDA dependency.
#include <algorithm>
#include <cstdint>

#define SXBH_API extern "C" __attribute__((visibility("default")))

// atoms: int32 [max_atoms, 6] = (seq_index, q_start_idx, q_len, global_q_idx, kv_blocks, total_extent)
// returns the number of atoms written, or -1 when max_atoms is too small.
SXBH_API int sxbh_build_atoms(int32_t* atoms, int max_atoms, const int32_t* seq_start, const int32_t* seq_tokens,
                              const int32_t* seq_seen, int n_seqs, int q_block, int kv_block) {
  int n = 0;
  for (int i = 0; i < n_seqs; ++i) {
    int cur = seq_start[i], global = seq_seen[i], remaining = seq_tokens[i];
    while (remaining > 0) {
      if (n >= max_atoms) return -1;
      const int q_len = std::min(remaining, q_block);
      const int end = global + q_len;
      int32_t* a = atoms + (long long)n * 6;
      a[0] = i;
      a[1] = cur;
      a[2] = q_len;
      a[3] = global;
      a[4] = (end + kv_block - 1) / kv_block;
      a[5] = end;
      cur += q_len;
      global += q_len;
      remaining -= q_len;
      ++n;
    }
  }
  return n;
}

// tok_seq / tok_pos: int32 [total tokens]; last_tok: int32 [n_seqs]; returns total tokens.
SXBH_API int sxbh_fill_ragged_batch(int32_t* tok_seq, int32_t* tok_pos, int32_t* last_tok, const int32_t* seq_tokens,
                                    const int32_t* seq_seen, int n_seqs) {
  int t = 0;
  for (int i = 0; i < n_seqs; ++i) {
    for (int j = 0; j < seq_tokens[i]; ++j, ++t) {
      tok_seq[t] = i;
      tok_pos[t] = seq_seen[i] + j;
    }
    last_tok[i] = t - 1;
  }
  return t;
}
