// Issue-rate micro-benchmark for the instruction mix of the node phase (sm_100a): which pipe each op uses and
// whether pairs of ops overlap.  One block of 128 threads per SM slot x 8; each test is a long unrolled chain
// of independent ops (8 accumulators) so that latency is hidden and the issue rate shows.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
template <int MODE>
__global__ void __launch_bounds__(128) k(uint32_t* out, uint32_t seed) {
  uint32_t a[8]; float f[8]; float2 g[8];
  for (int i = 0; i < 8; i++) { a[i] = seed + threadIdx.x * 8 + i; f[i] = 1.0f + i + seed; g[i] = make_float2(f[i], f[i] + 1.f); }
  const uint32_t magic = 0x3F800000u | (seed >> 31);
  const float s = __uint_as_float(0x3F800001u + seed), o = 0.5f + seed;
  const float2 s2 = make_float2(s, s), o2 = make_float2(o, o);
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) a[i] = __byte_perm(a[i], magic, 0x7614u);                       // PRMT
      if (MODE == 1) a[i] = __dp4a(a[i], 0x00004000u, magic);                  // IDP.4A
      if (MODE == 2) f[i] = fmaf(f[i], s, o);                                        // FFMA
      if (MODE == 3) g[i] = __ffma2_rn(g[i], s2, o2);                                // FFMA2
      if (MODE == 4) f[i] = fmaxf(fmaxf(f[i], s), o);                                // FMNMX3?
      if (MODE == 5) { a[i] = __byte_perm(a[i], magic, 0x7614u); f[i] = fmaf(f[i], s, o); }       // PRMT + FFMA
      if (MODE == 6) { a[i] = __dp4a(a[i], 0x00004000u, magic); f[i] = fmaf(f[i], s, o); }  // IDP + FFMA
      if (MODE == 7) { a[i] = __dp4a(a[i], 0x00004000u, magic); g[i] = __ffma2_rn(g[i], s2, o2); }  // IDP + FFMA2
      if (MODE == 8) { a[i] = __byte_perm(a[i], magic, 0x7614u); g[i] = __ffma2_rn(g[i], s2, o2); }      // PRMT + FFMA2
      if (MODE == 9) { a[i] = __dp4a(a[i], 0x00004000u, magic); a[i] = __byte_perm(a[i], magic, 0x7614u); }  // IDP + PRMT
      if (MODE == 10) f[i] = f[i] + s;                                               // FADD
      if (MODE == 11) { f[i] = f[i] + s; a[i] = __byte_perm(a[i], magic, 0x7614u); } // FADD + PRMT
      if (MODE == 12) a[i] = a[i] * 0x101u + seed;                                   // IMAD
      if (MODE == 13) { a[i] = __dp4a(a[i], 0x00004000u, magic); f[i] = fmaf(f[i], s, o); f[i] = fmaxf(fmaxf(f[i], s), o);}  // IDP+FFMA+FMNMX3
      if (MODE == 14) { a[i] = __funnelshift_l(a[i], a[(i + 1) & 7], 1); }          // SHF
      if (MODE == 15) { a[i] = (a[i] | a[(i + 1) & 7] | a[(i + 2) & 7]); }          // LOP3
    }
  }
  uint32_t r = 0;
  for (int i = 0; i < 8; i++) r ^= a[i] ^ __float_as_uint(f[i]) ^ __float_as_uint(g[i].x) ^ __float_as_uint(g[i].y);
  if (r == 0x12345678u) out[0] = r;
}

template <int MODE>
void run(const char* name, int ops_per_iter, uint32_t* d) {
  int sms = 148, bps = 8;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<sms * bps, 128>>>(d, 0); cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<sms * bps, 128>>>(d, 0);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  // warp-instructions issued per SMSP per cycle (assuming 1965 MHz)
  double winst = (double)sms * bps * 4 * ITERS * 8 * ops_per_iter;
  double cyc = ms * 1e-3 * 1.965e9;
  printf("%-22s %8.3f ms  %.3f warp-inst/clk/SMSP\n", name, ms, winst / (cyc * sms * 4));
}

int main() {
  uint32_t* d; cudaMalloc(&d, 4);
  run<0>("PRMT", 1, d); run<1>("IDP.4A", 1, d); run<2>("FFMA", 1, d); run<3>("FFMA2", 1, d); run<4>("FMNMX3", 1, d);
  run<10>("FADD", 1, d); run<12>("IMAD", 1, d); run<14>("SHF", 1, d); run<15>("LOP3", 1, d);
  run<5>("PRMT+FFMA", 2, d); run<6>("IDP+FFMA", 2, d); run<7>("IDP+FFMA2", 2, d); run<8>("PRMT+FFMA2", 2, d);
  run<9>("IDP+PRMT", 2, d); run<11>("FADD+PRMT", 2, d); run<13>("IDP+FFMA+FMNMX3", 3, d);
  return 0;
}
