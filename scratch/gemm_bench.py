import sys, torch
sys.path.insert(0, '.')
from llm_sharding_demo_b200 import ops
def bench(M, N, K, out_bf16=True, gelu=False):
    A = torch.randn(M, K, device="cuda").bfloat16(); W = (torch.randn(N, K, device="cuda") * 0.02).bfloat16()
    bias = torch.randn(N, device="cuda")
    for _ in range(3): ops.gemm(A, W, bias=bias, out_bf16=out_bf16, gelu=gelu)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): ops.gemm(A, W, bias=bias, out_bf16=out_bf16, gelu=gelu)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    # torch reference (cuBLAS) for context only
    for _ in range(3): torch.matmul(A, W.t())
    e0.record()
    for _ in range(10): torch.matmul(A, W.t())
    e1.record(); torch.cuda.synchronize()
    ms_ref = e0.elapsed_time(e1) / 10
    print(f"M={M} N={N} K={K}: {ms:.3f} ms = {2*M*N*K/ms/1e9:.1f} TFLOP/s   (cuBLAS {ms_ref:.3f} ms = {2*M*N*K/ms_ref/1e9:.1f})", flush=True)
for shp in [(16384, 2304, 768), (16384, 3072, 768), (16384, 768, 3072), (16384, 768, 768), (4096, 4800, 1600), (28672, 6400, 1600)]:
    bench(*shp)
