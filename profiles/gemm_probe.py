"""Runs the tcgen05 GEMM (gp_op_gemm through the C ABI) on two prefill-sized shapes; used by
profiles/run_ncu.sh for the ncu capture and prints CUDA-event TFLOP/s when run plainly."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_sharding_demo_b200 import ops  # noqa: E402

for (M, N, K) in [(16384, 3072, 768), (8192, 8192, 8192)]:
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") * 0.02).bfloat16()
    bias = torch.randn(N, device="cuda")
    for _ in range(3):
        ops.gemm(A, W, bias=bias, out_bf16=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.gemm(A, W, bias=bias, out_bf16=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"gemm_tc M={M} N={N} K={K}: {ms:.3f} ms = {2 * M * N * K / ms / 1e9:.1f} TFLOP/s")
