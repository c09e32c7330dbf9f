"""TEST INFRASTRUCTURE ONLY (oracle): CPU restatements of the two kernels either side of the decode path.

advance_step_flashattn  /root/reference/csrc/prepare_inputs/advance_step.cu:13-65
merge_attn_states       /root/reference/csrc/attention/merge_attn_states.cu:12-89; the reference's own torch golden is
                        tests/kernels/attention/test_merge_attn_states.py:16-56 (`merge_attn_states_torch`), which this
                        follows (section 2.2 of arXiv 2501.01005).
Parity: advance_step is integer work -> bit-exact.  merge_attn_states: fp32 math, one rounding to the output dtype
-> 1e-2 rel / 2e-2 abs bar (observed: <= 1 ulp of the output dtype).  Parity unpinned for advance_step beyond the
source restatement (the reference holds no golden vectors for it).
"""
import torch


def advance_step_flashattn(num_seqs, num_queries, block_size, input_tokens, sampled_token_ids, input_positions,
                           seq_lens, slot_mapping, block_tables):
    """In place on CPU tensors, a plain loop over the queries (advance_step.cu:36-64) + the padding rows (:21-30)."""
    for i in range(num_queries, num_seqs):
        input_tokens[i] = 0
        input_positions[i] = 0
        slot_mapping[i] = -1
    for i in range(num_queries):
        input_tokens[i] = sampled_token_ids[i, 0]
        nxt = int(seq_lens[i]) + 1
        pos = nxt - 1
        seq_lens[i] = nxt
        input_positions[i] = pos
        slot_mapping[i] = int(block_tables[i, pos // block_size]) * block_size + pos % block_size


def merge_attn_states(prefix_output, prefix_lse, suffix_output, suffix_lse):
    """-> (output [T, H, D] in prefix_output.dtype, output_lse fp32 [H, T]).  +-inf in an lse counts as -inf
    (merge_attn_states.cu:45-46; test_merge_attn_states.py:27-30)."""
    p_lse = prefix_lse.float().clone()
    s_lse = suffix_lse.float().clone()
    p_lse[torch.isinf(p_lse)] = float("-inf")
    s_lse[torch.isinf(s_lse)] = float("-inf")
    m = torch.maximum(p_lse, s_lse)
    p_se, s_se = torch.exp(p_lse - m), torch.exp(s_lse - m)
    out_se = p_se + s_se
    p_scale = (p_se / out_se).t().unsqueeze(-1)     # [T, H, 1]
    s_scale = (s_se / out_se).t().unsqueeze(-1)
    out = (prefix_output.float() * p_scale + suffix_output.float() * s_scale).to(prefix_output.dtype)
    return out, torch.log(out_se) + m
