// umma_probe.cu - standalone sm_100a probe of the tcgen05 operand conventions the backward kernel relies on:
//   (1) dgrad: A = G K-major (128 x Kn), B = forward weight buffer (rows n, K-major over k) read MN-major
//   (2) wgrad: A = G buffer read MN-major (M = 64 features, K = 128 samples), B = X buffer read MN-major,
//       accumulator with M = 64 -> which TMEM lanes hold which rows
//   (3) mixed operand types (A bf16, B fp16) inside kind::f16
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../nerf_emitter_b200/csrc umma_probe.cu -o umma_probe
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_bf16.h>
#include "tc_common.cuh"

using namespace nemit;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

// generic single-MMA-chain kernel: smem image of A and B given by the host, descriptors given by the host
struct ProbeArgs {
  const uint8_t* a_img; int a_bytes;
  const uint8_t* b_img; int b_bytes;
  uint32_t a_lbo, a_sbo, a_step;   // byte offsets; a_step = start-address advance per k-step
  uint32_t b_lbo, b_sbo, b_step;
  uint32_t idesc;
  int ksteps;
  int ncols;          // columns to read back (multiple of 16)
  float* out;         // (128 lanes, ncols)
};

__global__ void __launch_bounds__(128) k_probe(ProbeArgs P) {
  extern __shared__ __align__(128) unsigned char smem[];
  uint8_t* a = smem;
  uint8_t* b = smem + 65536;
  __shared__ uint64_t mbar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, wid = tid >> 5;
  for (int i = tid; i < P.a_bytes; i += 128) a[i] = P.a_img[i];
  for (int i = tid; i < P.b_bytes; i += 128) b[i] = P.b_img[i];
  if (tid == 0) { tc::mbar_init(&mbar, 1); tc::fence_mbar_init(); }
  if (wid == 0) tc::tmem_alloc<128>(&slot);
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = slot;
  // clear the accumulator region first with a zero MMA? simply rely on accumulate = 0 on the first k-step
  if (tid == 0) {
    const uint32_t a0 = tc::smem_u32(a), b0 = tc::smem_u32(b);
    for (int j = 0; j < P.ksteps; ++j)
      tc::mma_f16_ss(tmem, tc::smem_desc(a0 + j * P.a_step, P.a_lbo, P.a_sbo), tc::smem_desc(b0 + j * P.b_step, P.b_lbo, P.b_sbo),
                     P.idesc, j > 0 ? 1u : 0u);
    tc::mma_commit(&mbar);
  }
  tc::mbar_wait(&mbar, 0);
  tc::tc_fence_after();
  const uint32_t taddr = tmem + ((uint32_t)(wid * 32) << 16);
  for (int c = 0; c < P.ncols; c += 16) {
    float v[16];
    tc::tmem_ld16(taddr + c, v);
    tc::tmem_ld_wait();
    for (int i = 0; i < 16; ++i) P.out[tid * P.ncols + c + i] = v[i];
  }
  tc::tc_fence_before();
  __syncthreads();
  if (wid == 0) tc::tmem_dealloc<128>(tmem);
}

static uint16_t f2h(float f) { __half h = __float2half_rn(f); return *reinterpret_cast<uint16_t*>(&h); }
static uint16_t f2bf(float f) { __nv_bfloat16 h = __float2bfloat16_rn(f); return *reinterpret_cast<uint16_t*>(&h); }
static float h2f(uint16_t u) { __half h = *reinterpret_cast<__half*>(&u); return __half2float(h); }
static float bf2f(uint16_t u) { __nv_bfloat16 h = *reinterpret_cast<__nv_bfloat16*>(&u); return __bfloat162float(h); }

// the repo's activation / weight buffer layout: (row r, col k) of an R-row buffer
static size_t off_rk(int R, int r, int k) { return (size_t)(k / 8) * R * 16 + (size_t)(r / 8) * 128 + (r % 8) * 16 + (k % 8) * 2; }

static uint32_t idesc(int M, int N, int a_bf16, int b_bf16, int a_mn, int b_mn) {
  return (1u << 4) | ((uint32_t)a_bf16 << 7) | ((uint32_t)b_bf16 << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

static std::vector<float> run(const std::vector<uint8_t>& a, const std::vector<uint8_t>& b, ProbeArgs P) {
  uint8_t *da, *db; float* dout;
  CK(cudaMalloc(&da, a.size())); CK(cudaMalloc(&db, b.size())); CK(cudaMalloc(&dout, 128 * P.ncols * 4));
  CK(cudaMemcpy(da, a.data(), a.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, b.data(), b.size(), cudaMemcpyHostToDevice));
  P.a_img = da; P.a_bytes = (int)a.size(); P.b_img = db; P.b_bytes = (int)b.size(); P.out = dout;
  CK(cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
  k_probe<<<1, 128, 131072>>>(P);
  CK(cudaDeviceSynchronize());
  std::vector<float> out(128 * P.ncols);
  CK(cudaMemcpy(out.data(), dout, out.size() * 4, cudaMemcpyDeviceToHost));
  cudaFree(da); cudaFree(db); cudaFree(dout);
  return out;
}

static float frand() { return (float)rand() / RAND_MAX * 2.f - 1.f; }

int main(int argc, char** argv) {
  srand(1);
  const int test = argc > 1 ? atoi(argv[1]) : 1;
  // ---------------- (1) dgrad: D[m][k] = sum_n G[m][n] W[n][k]; G 128 x 64 (K-major A), W buffer 64 rows(n) x 32 cols(k)
  if (test == 1) { const int mixed = 0;
    const int NN = 64, KK = 32;
    std::vector<float> G(128 * NN), W(NN * KK);
    std::vector<uint8_t> a(128 * NN * 2), b(NN * KK * 2);
    for (int m = 0; m < 128; ++m) for (int n = 0; n < NN; ++n) {
      uint16_t u = mixed ? f2bf(frand()) : f2h(frand());
      G[m * NN + n] = mixed ? bf2f(u) : h2f(u);
      *reinterpret_cast<uint16_t*>(&a[off_rk(128, m, n)]) = u;
    }
    for (int n = 0; n < NN; ++n) for (int k = 0; k < KK; ++k) {
      uint16_t u = f2h(frand()); W[n * KK + k] = h2f(u);
      *reinterpret_cast<uint16_t*>(&b[off_rk(NN, n, k)]) = u;
    }
    ProbeArgs P{};
    P.a_lbo = 128 * 16; P.a_sbo = 128; P.a_step = 2 * 128 * 16;       // K-major A, k-step = 2 K-chunks
    // B read MN-major: MN = k (8 contiguous), K = n rows (16 B apart); MN groups at NN*16, K groups at 128
    P.b_lbo = 128; P.b_sbo = NN * 16; P.b_step = 2 * 128;              // 16 n-rows per k-step = 2 K-groups
    P.idesc = idesc(128, KK, mixed, 0, 0, 1);
    P.ksteps = NN / 16; P.ncols = 32;
    auto out = run(a, b, P);
    double err = 0;
    for (int m = 0; m < 128; ++m) for (int k = 0; k < KK; ++k) {
      double r = 0; for (int n = 0; n < NN; ++n) r += (double)G[m * NN + n] * W[n * KK + k];
      err = fmax(err, fabs(r - out[m * 32 + k]));
    }
    printf("[1] dgrad (A K-major %s, B fwd-weight buffer MN-major fp16): max abs err %.3e\n", mixed ? "bf16" : "fp16", err);
    if (err > 1e-2) {
      // try swapped lbo/sbo
      std::swap(P.b_lbo, P.b_sbo);
      out = run(a, b, P);
      err = 0;
      for (int m = 0; m < 128; ++m) for (int k = 0; k < KK; ++k) {
        double r = 0; for (int n = 0; n < NN; ++n) r += (double)G[m * NN + n] * W[n * KK + k];
        err = fmax(err, fabs(r - out[m * 32 + k]));
      }
      printf("    with LBO/SBO swapped: max abs err %.3e\n", err);
    }
  }
  // ---------------- (2) wgrad: D[n][k] = sum_m G[m][n] X[m][k]; G 128 x 64 buffer, X 128 x KX buffer, M = 64
  if (test == 2) for (int KX : {64, 32, 16, 8}) { const int mixed = 0;
    const int NN = 64;
    std::vector<float> G(128 * NN), X(128 * KX);
    std::vector<uint8_t> a(128 * NN * 2), b(128 * KX * 2);
    for (int m = 0; m < 128; ++m) for (int n = 0; n < NN; ++n) {
      uint16_t u = mixed ? f2bf(frand()) : f2h(frand());
      G[m * NN + n] = mixed ? bf2f(u) : h2f(u);
      *reinterpret_cast<uint16_t*>(&a[off_rk(128, m, n)]) = u;
    }
    for (int m = 0; m < 128; ++m) for (int k = 0; k < KX; ++k) {
      uint16_t u = f2h(frand()); X[m * KX + k] = h2f(u);
      *reinterpret_cast<uint16_t*>(&b[off_rk(128, m, k)]) = u;
    }
    ProbeArgs P{};
    // both MN-major: MN groups (features / 8) at 128*16 bytes, K groups (samples / 8) at 128 bytes
    P.a_lbo = 128; P.a_sbo = 128 * 16; P.a_step = 2 * 128;
    P.b_lbo = 128; P.b_sbo = 128 * 16; P.b_step = 2 * 128;
    P.idesc = idesc(64, KX, mixed, 0, 1, 1);
    P.ksteps = 128 / 16; P.ncols = 64;
    auto out = run(a, b, P);
    std::vector<double> ref(NN * KX);
    for (int n = 0; n < NN; ++n) for (int k = 0; k < KX; ++k) {
      double r = 0; for (int m = 0; m < 128; ++m) r += (double)G[m * NN + n] * X[m * KX + k];
      ref[n * KX + k] = r;
    }
    // hypothesis: row n lives in lane (n / 16) * 32 + n % 16
    double err = 0;
    for (int n = 0; n < NN; ++n) for (int k = 0; k < KX; ++k)
      err = fmax(err, fabs(ref[n * KX + k] - out[((n / 16) * 32 + n % 16) * 64 + k]));
    printf("[2] wgrad M=64 N=%d (%s A): max abs err with lane = (n/16)*32 + n%%16: %.3e\n", KX, mixed ? "bf16" : "fp16", err);
    if (err > 1e-2) {
      std::vector<uint8_t> dummy;
      printf("    lane -> row map (by best match): ");
      for (int lane = 0; lane < 128; ++lane) {
        int best = -1; double be = 1e30;
        for (int n = 0; n < NN; ++n) {
          double e = 0; for (int k = 0; k < KX; ++k) e = fmax(e, fabs(ref[n * KX + k] - out[lane * 64 + k]));
          if (e < be) { be = e; best = n; }
        }
        if (be < 1e-2) printf("%d:%d ", lane, best);
      }
      printf("\n");
    }
  }
  // ---------------- (3) row sums with a ones operand: D[n][j] = sum_m G[m][n] * 1, B = ones (8 x 128 MN-major)
  if (test == 3 || test == 4) {
    const int NN = 64;
    std::vector<float> G(128 * NN);
    std::vector<uint8_t> a(128 * NN * 2), b(128 * 8 * 2);
    for (int m = 0; m < 128; ++m) for (int n = 0; n < NN; ++n) {
      uint16_t u = f2h(frand()); G[m * NN + n] = h2f(u);
      *reinterpret_cast<uint16_t*>(&a[off_rk(128, m, n)]) = u;
    }
    for (int i = 0; i < 128 * 8; ++i) *reinterpret_cast<uint16_t*>(&b[i * 2]) = f2h(1.f);
    ProbeArgs P{};
    P.a_lbo = 128; P.a_sbo = 128 * 16; P.a_step = 2 * 128;
    P.b_lbo = 128; P.b_sbo = 128 * 16; P.b_step = 2 * 128;
    P.idesc = idesc(64, 8, 0, 0, 1, 1);
    if (test == 4) { P.b_lbo = 0; P.b_step = 0; }   // zero-stride ones: one 128-byte core matrix reused
    P.ksteps = 8; P.ncols = 16;
    auto out = run(a, b, P);
    double err = 0;
    for (int n = 0; n < NN; ++n) {
      double r = 0; for (int m = 0; m < 128; ++m) r += G[m * NN + n];
      err = fmax(err, fabs(r - out[((n / 16) * 32 + n % 16) * 16 + 0]));
    }
    printf("[3] column sums via ones operand (N = 8): max abs err %.3e\n", err);
  }
  // ---------------- (5) column sums with ones as the A operand (M = 64, SBO = 0): D[i][n] = sum_m G[m][n]
  if (test == 5) {
    const int NN = 16;
    std::vector<float> G(128 * NN);
    std::vector<uint8_t> a(128 * 8 * 2), b(128 * NN * 2);
    for (int m = 0; m < 128; ++m) for (int n = 0; n < NN; ++n) {
      uint16_t u = f2h(frand()); G[m * NN + n] = h2f(u);
      *reinterpret_cast<uint16_t*>(&b[off_rk(128, m, n)]) = u;
    }
    for (int i = 0; i < 128 * 8; ++i) *reinterpret_cast<uint16_t*>(&a[i * 2]) = f2h(1.f);
    ProbeArgs P{};
    P.a_lbo = 128; P.a_sbo = 0; P.a_step = 2 * 128;
    P.b_lbo = 128; P.b_sbo = 128 * 16; P.b_step = 2 * 128;
    P.idesc = idesc(64, NN, 0, 0, 1, 1);
    P.ksteps = 8; P.ncols = 16;
    auto out = run(a, b, P);
    double err = 0;
    for (int i = 0; i < 64; ++i) for (int n = 0; n < NN; ++n) {
      double r = 0; for (int m = 0; m < 128; ++m) r += G[m * NN + n];
      err = fmax(err, fabs(r - out[((i / 16) * 32 + i % 16) * 16 + n]));
    }
    printf("[5] column sums with ones as A (M = 64, SBO = 0, N = 16): max abs err %.3e\n", err);
  }
  return 0;
}
