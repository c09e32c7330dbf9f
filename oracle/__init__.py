"""ORACLE -- TEST INFRASTRUCTURE ONLY.

CPU restatement (torch CPU tensors + numpy, plain Python loops for small cases) of the reference's
algorithms for the decode hot path, each function citing the /root/reference file:line it follows.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this package; the product (vllm_b200/) never does.

Parity pinning (see DESIGN.md "Oracle"):
  * paged attention (bf16/f16, block 16) and reshape_and_cache are checked against the reference's
    OWN CPU kernels compiled from /root/reference/csrc/cpu/{attention,cache}.cpp
    (oracle/_ref/libvllm_ref_cpu.so, recipe oracle/build_ref.py) -- tests/test_oracle_pinning.py;
  * golden vectors produced by those reference kernels are committed under tests/golden/
    (generator: tests/golden/make_golden.py);
  * moe_align_block_size is checked against the worked example in the reference docstring
    (vllm/model_executor/layers/fused_moe/moe_align_block_size.py:52-66);
  * on the GPU box every op is additionally compared with the reference's own CUDA kernels
    (oracle/_ref/libvllm_ref_cuda.so, torch.ops._ref_C.*) and, where loadable, the vllm wheel.
"""
