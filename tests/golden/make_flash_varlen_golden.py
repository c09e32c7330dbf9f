"""Generates tests/golden/flash_varlen_*.npz: outputs of the REFERENCE test suite's own golden function for
flash_attn_varlen_func with a paged KV cache (`ref_paged_attn`,
/root/reference/tests/kernels/attention/test_flash_attn.py:25-77) on small seeded inputs.  The function is taken from
the reference file as it lies (its source is parsed and executed here, nothing is copied into the repository); the
module itself cannot be imported without a GPU build of vllm.  Run in the build container only:

    python tests/golden/make_flash_varlen_golden.py
"""
import ast
import os
from typing import Optional  # noqa: F401  (used by the reference function's annotations)

import numpy as np
import torch

REF = "/root/reference/tests/kernels/attention/test_flash_attn.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_ref_paged_attn():
    tree = ast.parse(open(REF).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "ref_paged_attn")
    ns = {"torch": torch, "Optional": Optional}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), REF, "exec"), ns)
    return ns["ref_paged_attn"]


def main():
    ref = load_ref_paged_attn()
    cases = {
        # the reference test's first seq_lens list (:192-193) at one of its head configurations
        "flash_varlen_ref_case": dict(seq_lens=[(1, 1328), (5, 18), (129, 463)], H=8, Hkv=2, D=128, bs=16, nb=32, seed=0),
        "flash_varlen_prefill_gqa7": dict(seq_lens=[(70, 70), (33, 200), (1, 97)], H=14, Hkv=2, D=64, bs=32, nb=16, seed=1),
        # the reference test's sliding_window / soft_cap parameters (:86-88, :201-203)
        "flash_varlen_window256": dict(seq_lens=[(1, 523), (130, 400), (5, 18)], H=4, Hkv=2, D=64, bs=16, nb=24, seed=2,
                                       sliding_window=256),
        "flash_varlen_softcap10_window": dict(seq_lens=[(1, 300), (90, 333)], H=4, Hkv=4, D=64, bs=16, nb=24, seed=3,
                                              sliding_window=100, soft_cap=10.0),
        "flash_varlen_softcap50": dict(seq_lens=[(40, 200), (1, 77)], H=8, Hkv=2, D=64, bs=16, nb=16, seed=4, soft_cap=50.0),
    }
    for name, c in cases.items():
        torch.manual_seed(c["seed"])
        ql = [x[0] for x in c["seq_lens"]]
        kl = [x[1] for x in c["seq_lens"]]
        q = torch.randn(sum(ql), c["H"], c["D"], dtype=torch.bfloat16)
        kc = torch.randn(c["nb"], c["bs"], c["Hkv"], c["D"], dtype=torch.bfloat16)
        vc = torch.randn_like(kc)
        maxb = (max(kl) + c["bs"] - 1) // c["bs"]
        bt = torch.randint(0, c["nb"], (len(ql), maxb), dtype=torch.int32)
        scale = c["D"] ** -0.5
        out = ref(query=q.clone(), key_cache=kc, value_cache=vc, query_lens=ql, kv_lens=kl, block_tables=bt, scale=scale,
                  sliding_window=c.get("sliding_window"), soft_cap=c.get("soft_cap"))
        as_i16 = lambda t: t.contiguous().view(torch.int16).numpy()  # noqa: E731
        np.savez_compressed(os.path.join(HERE, name + ".npz"), q=as_i16(q), key_cache=as_i16(kc), value_cache=as_i16(vc),
                            block_tables=bt.numpy(), query_lens=np.array(ql), kv_lens=np.array(kl),
                            scale=np.float32(scale), out=as_i16(out.to(torch.bfloat16)),
                            sliding_window=np.int64(c.get("sliding_window") or -1),
                            soft_cap=np.float32(c.get("soft_cap") or 0.0))
        print(name, tuple(out.shape), os.path.getsize(os.path.join(HERE, name + ".npz")))


if __name__ == "__main__":
    main()
