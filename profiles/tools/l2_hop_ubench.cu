// Microbenchmark behind DESIGN.md section 6: dependent HMMA latency, LDS->HMMA step cost, and the L2 store->poll hop between two SMs.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench l2_hop_ubench.cu ; measured on B200: 830 cycles per hop.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__global__ void k(long long* out, int n, int warps_active) {
  __shared__ uint32_t sm[4096];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = i;
  __syncthreads();
  if ((threadIdx.x >> 5) >= warps_active) return;
  float d[4] = {0, 0, 0, 0}, e[4] = {0, 0, 0, 0};
  uint32_t a = threadIdx.x, b = 3;
  // dependent HMMA chain
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) mma16816(d, a, a, a, a, b, b);
  long long t1 = clock64();
  // two independent chains
  for (int i = 0; i < n; ++i) { mma16816(d, a, a, a, a, b, b); mma16816(e, a, a, a, a, b, b); }
  long long t2 = clock64();
  // LDS.64 -> HMMA dependent chain (address depends on previous result low bits)
  uint32_t addr = (uint32_t)__cvta_generic_to_shared(sm) + (threadIdx.x & 31) * 8;
  uint32_t x0 = 0, x1 = 0;
  long long t3 = clock64();
  for (int i = 0; i < n; ++i) {
    asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];" : "=r"(x0), "=r"(x1) : "r"(addr + (x0 & 0)));
  }
  long long t4 = clock64();
  for (int i = 0; i < n; ++i) {
    asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];" : "=r"(x0), "=r"(x1) : "r"(addr + i * 256));
    mma16816(d, x0, x1, x0, x1, x0, x1);
  }
  long long t5 = clock64();
  if (threadIdx.x == 0) {
    out[0] = t1 - t0; out[1] = t2 - t1; out[2] = t4 - t3; out[3] = t5 - t4;
  }
  if (d[0] + e[0] + x0 == 12345.f) out[7] = 1;
}
__global__ void pingpong(unsigned long long* buf, long long* out, int n) {
  // block 0 and block 1 bounce a word: measures store->visible->load round trip between two SMs through L2
  unsigned long long* p = buf + 0, *q = buf + 128;
  if (threadIdx.x != 0) return;
  long long t0 = clock64();
  for (int i = 1; i <= n; ++i) {
    if (blockIdx.x == 0) {
      asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"((unsigned long long)i) : "memory");
      unsigned long long v;
      do { asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(q) : "memory"); } while (v != (unsigned long long)i);
    } else {
      unsigned long long v;
      do { asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); } while (v != (unsigned long long)i);
      asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(q), "l"((unsigned long long)i) : "memory");
    }
  }
  long long t1 = clock64();
  if (blockIdx.x == 0) out[4] = t1 - t0;
}
int main() {
  long long* out; cudaMallocManaged(&out, 64);
  unsigned long long* buf; cudaMalloc(&buf, 4096); cudaMemset(buf, 0, 4096);
  for (int wa : {1, 4, 15}) {
    for (int rep = 0; rep < 2; ++rep) { k<<<1, 512>>>(out, 64, wa); cudaDeviceSynchronize(); }
    printf("warps=%d n=64: dep HMMA %.1f cyc/op, 2 indep chains %.1f cyc/pair, dep LDS.64 %.1f cyc, LDS->HMMA %.1f cyc/step\n", wa,
           out[0] / 64.0, out[1] / 64.0, out[2] / 64.0, out[3] / 64.0);
  }
  for (int rep = 0; rep < 2; ++rep) { cudaMemset(buf, 0, 4096); pingpong<<<2, 32>>>(buf, out, 1000); cudaDeviceSynchronize(); }
  printf("L2 ping-pong: %.1f cyc per round trip (2 hops)\n", out[4] / 1000.0);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
