// rowbench.cu -- what one Gauss-Seidel normal-row update costs a warp that runs alone.
// Synthetic entries in shared memory (unit rows, W = 14), 8 entries x 50 iterations, variants:
//   0  load -> gap -> clamp -> axpy, one entry at a time (the round-1 loop)
//   1  two register buffers, the next entry loaded during the update (solve_warp_kernel now)
//   2  look-ahead: gap_j = (b_j - u_j . y_after(j-2)) - (u_j . u_(j-1)) d_(j-1); the long dot and
//      the previous axpy run in the shadow of the three-instruction chain
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o rowbench rowbench.cu && ./rowbench
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

constexpr int W = 14, ES = 64, NE = 8, ITERS = 50;

__device__ __forceinline__ void load16(const float *e, float (&r)[16]) {
#pragma unroll
  for (int i = 0; i < 16; i += 4) {
    const float4 v = *reinterpret_cast<const float4 *>(e + i);
    r[i] = v.x; r[i + 1] = v.y; r[i + 2] = v.z; r[i + 3] = v.w;
  }
}
__device__ __forceinline__ float gap4(const float (&r)[16], const float (&y)[W], const float extra) {
  float a0 = r[W] + extra, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
  for (int a = 0; a < W; a++) {
    if ((a & 3) == 0) a0 = fmaf(-r[a], y[a], a0);
    else if ((a & 3) == 1) a1 = fmaf(-r[a], y[a], a1);
    else if ((a & 3) == 2) a2 = fmaf(-r[a], y[a], a2);
    else a3 = fmaf(-r[a], y[a], a3);
  }
  return (a0 + a1) + (a2 + a3);
}

template <int V>
__global__ void __launch_bounds__(32) rows(const float *init, float *out, long long *cycles, const int n_act_in) {
  __shared__ __align__(16) float ent[NE + 1][ES];
  for (int i = threadIdx.x; i < (NE + 1) * ES; i += 32) (&ent[0][0])[i] = init[i];
  __syncwarp();
  float y[W];
#pragma unroll
  for (int a = 0; a < W; a++) y[a] = init[(NE + 1) * ES - 1 - a] * (float)(threadIdx.x >> 5);
  float res = 0.f, budget = 1e30f;
  const int n_act = n_act_in;
  const long long t0 = clock64();
  for (int it = 0; it < ITERS; it++) {
    if (V == 0) {
      for (int k = 0; k < n_act; k++) {
        float r[16];
        load16(&ent[k][0], r);
        const float lam = ent[k][16];
        const float sum = fmaxf(gap4(r, y, lam), 0.f);
        const float d = sum - lam;
        ent[k][16] = sum;
#pragma unroll
        for (int a = 0; a < W; a++) y[a] = fmaf(r[a], d, y[a]);
        const float rr = d * r[W + 1];
        res = fmaxf(res, rr * rr);
        budget -= fabsf(d);
      }
    } else if (V == 1) {
      float A[16], Bf[16], lamA, lamB;
      load16(&ent[0][0], A); lamA = ent[0][16];
      for (int k = 0;;) {
        load16(&ent[k + 1][0], Bf); lamB = ent[k + 1][16];
        {
          const float sum = fmaxf(gap4(A, y, lamA), 0.f), d = sum - lamA;
          ent[k][16] = sum;
#pragma unroll
          for (int a = 0; a < W; a++) y[a] = fmaf(A[a], d, y[a]);
          const float rr = d * A[W + 1]; res = fmaxf(res, rr * rr); budget -= fabsf(d);
        }
        if (++k >= n_act) break;
        load16(&ent[k + 1][0], A); lamA = ent[k + 1][16];
        {
          const float sum = fmaxf(gap4(Bf, y, lamB), 0.f), d = sum - lamB;
          ent[k][16] = sum;
#pragma unroll
          for (int a = 0; a < W; a++) y[a] = fmaf(Bf[a], d, y[a]);
          const float rr = d * Bf[W + 1]; res = fmaxf(res, rr * rr); budget -= fabsf(d);
        }
        if (++k >= n_act) break;
      }
    } else if (V == 3 || V == 4 || V == 5) {
      // V1 + the real loop's decision "screen the speculative rows?" on (range, budget), where
      // budget -= |d| hangs off the end of the chain.  3: decision first (what the compiler makes
      // of the product loop); 4: the row's own gap pinned in front of the decision; 5: 4 + the
      // decision taken on the budget as it was one row earlier (one row of slack in the bound)
      float A[16], Bf[16], lamA, lamB; unsigned rA, rB;
      load16(&ent[0][0], A); lamA = ent[0][16]; rA = __float_as_uint(ent[0][18]);
      float b_prev = budget;
      auto upd = [&](const int k, const float (&r)[16], const float lam, const unsigned range) -> bool {
        float g = 0.f;
        if (V >= 4) { g = gap4(r, y, lam); asm volatile("" : "+f"(g)); }
        const float bb = V == 5 ? b_prev : budget;
        if (range != 0 && bb <= 0.f) {
          // stand-in for the screen: a ballot and a reduction the compiler cannot drop
          const unsigned m = __ballot_sync(0xffffffffu, y[threadIdx.x & 7] > 0.f);
          budget = __uint_as_float(__reduce_min_sync(0xffffffffu, __float_as_uint(fabsf(y[0]) + 1.f)));
          if (m & range & 0x80000000u) return true;
        }
        if (V < 4) g = gap4(r, y, lam);
        const float sum = fmaxf(g, 0.f), d = sum - lam;
        ent[k][16] = sum;
#pragma unroll
        for (int a = 0; a < W; a++) y[a] = fmaf(r[a], d, y[a]);
        const float rr = d * r[W + 1]; res = fmaxf(res, rr * rr);
        b_prev = budget;
        budget -= fabsf(d);
        return false;
      };
      for (int k = 0;;) {
        load16(&ent[k + 1][0], Bf); lamB = ent[k + 1][16]; rB = __float_as_uint(ent[k + 1][18]);
        if (upd(k, A, lamA, rA)) break;
        if (++k >= n_act) break;
        load16(&ent[k + 1][0], A); lamA = ent[k + 1][16]; rA = __float_as_uint(ent[k + 1][18]);
        if (upd(k, Bf, lamB, rB)) break;
        if (++k >= n_act) break;
      }
    } else if (V == 6 || V == 7) {
      // the decision leaves the fast loop (forward branch, not taken on the common path); the
      // screen sits behind the loop and re-enters it.  7: four rows per trip instead of two.
      float A[16], Bf[16], lamA, lamB; unsigned rA, rB;
      bool screened = false;
      for (int k = 0;;) {
        bool need = false;
        load16(&ent[k][0], A); lamA = ent[k][16]; rA = __float_as_uint(ent[k][18]);
        auto upd = [&](const int kk, const float (&r)[16], const float lam) {
          const float sum = fmaxf(gap4(r, y, lam), 0.f), d = sum - lam;
          ent[kk][16] = sum;
#pragma unroll
          for (int a = 0; a < W; a++) y[a] = fmaf(r[a], d, y[a]);
          const float rr = d * r[W + 1]; res = fmaxf(res, rr * rr);
          budget -= fabsf(d);
        };
        for (;;) {
          if (rA != 0 && budget <= 0.f && !screened) { need = true; break; }
          screened = false;
          load16(&ent[k + 1][0], Bf); lamB = ent[k + 1][16]; rB = __float_as_uint(ent[k + 1][18]);
          upd(k, A, lamA);
          if (++k >= n_act) break;
          if (rB != 0 && budget <= 0.f) { need = true; break; }
          load16(&ent[k + 1][0], A); lamA = ent[k + 1][16]; rA = __float_as_uint(ent[k + 1][18]);
          upd(k, Bf, lamB);
          if (++k >= n_act) break;
          if (V == 7) {
            if (rA != 0 && budget <= 0.f) { need = true; break; }
            load16(&ent[k + 1][0], Bf); lamB = ent[k + 1][16]; rB = __float_as_uint(ent[k + 1][18]);
            upd(k, A, lamA);
            if (++k >= n_act) break;
            if (rB != 0 && budget <= 0.f) { need = true; break; }
            load16(&ent[k + 1][0], A); lamA = ent[k + 1][16]; rA = __float_as_uint(ent[k + 1][18]);
            upd(k, Bf, lamB);
            if (++k >= n_act) break;
          }
        }
        if (!need) break;
        const unsigned m = __ballot_sync(0xffffffffu, y[threadIdx.x & 7] > 0.f);
        budget = __uint_as_float(__reduce_min_sync(0xffffffffu, __float_as_uint(fabsf(y[0]) + 1.f)));
        screened = true;
        if (m & __float_as_uint(ent[k][18]) & 0x80000000u) break;
      }
    } else if (V == 8 || V == 9) {
      // look-ahead (variant 2) per iteration with the screen decision per row: the first row of
      // the pass takes the plain form, the pending row is flushed at the end of the pass
      float B0[16], B1[16], B2[16], l0, l1, l2, c0, c1, c2; unsigned r0, r1, r2;
      auto ld = [&](const int k, float (&b)[16], float &l, float &c, unsigned &r) {
        load16(&ent[k][0], b); l = ent[k][16]; c = ent[k][17]; r = __float_as_uint(ent[k][18]);
      };
      ld(0, B0, l0, c0, r0);
      ld(1, B1, l1, c1, r1);
      float g = gap4(B0, y, l0), dp = 0.f;   // row 0 against the fully applied y
      c0 = 0.f;
      bool stop = false;
      // step: row `cur` (gap g, coupling cc with the pending row `prev`), then prev into y and
      // the next row's gap against it
      auto step = [&](const int k, const float (&prev)[16], const float (&cur)[16], const float lc, const float cc,
                      const unsigned rc, const float (&nxt)[16], const float ln) {
        float gn = 0.f, sum, d;
        if (V == 9) {
          // everything that does not depend on the decision in front of it: the pending row into
          // y, the next row's gap, this row's clamp
#pragma unroll
          for (int a = 0; a < W; a++) y[a] = fmaf(prev[a], dp, y[a]);
          gn = gap4(nxt, y, ln);
          sum = fmaxf(fmaf(-cc, dp, g), 0.f); d = sum - lc;
          asm volatile("" : "+f"(gn), "+f"(d), "+f"(sum));
        }
        float gg = fmaf(-cc, dp, g);
        asm volatile("" : "+f"(gg));
        if (rc != 0 && budget <= 0.f) {
          const unsigned m = __ballot_sync(0xffffffffu, y[threadIdx.x & 7] > 0.f);
          budget = __uint_as_float(__reduce_min_sync(0xffffffffu, __float_as_uint(fabsf(y[0]) + 1.f)));
          if (m & rc & 0x80000000u) { stop = true; return; }
        }
        if (V != 9) {
          sum = fmaxf(gg, 0.f); d = sum - lc;
#pragma unroll
          for (int a = 0; a < W; a++) y[a] = fmaf(prev[a], dp, y[a]);
          gn = gap4(nxt, y, ln);
        }
        ent[k][16] = sum;
        g = gn;
        const float rr = d * cur[W + 1]; res = fmaxf(res, rr * rr); budget -= fabsf(d);
        dp = d;
      };
      float Z[16];
#pragma unroll
      for (int a = 0; a < 16; a++) Z[a] = 0.f;
      // rows 0, 1, 2 peeled by hand into a 3-buffer rotation
      step(0, Z, B0, l0, c0, r0, B1, l1);
      for (int k = 1; !stop;) {
        ld(k + 1, B2, l2, c2, r2);
        step(k, B0, B1, l1, c1, r1, B2, l2); if (stop || ++k >= n_act) { if (!stop) {
#pragma unroll
          for (int a = 0; a < W; a++) y[a] = fmaf(B1[a], dp, y[a]); } break; }
        ld(k + 1, B0, l0, c0, r0);
        step(k, B1, B2, l2, c2, r2, B0, l0); if (stop || ++k >= n_act) { if (!stop) {
#pragma unroll
          for (int a = 0; a < W; a++) y[a] = fmaf(B2[a], dp, y[a]); } break; }
        ld(k + 1, B1, l1, c1, r1);
        step(k, B2, B0, l0, c0, r0, B1, l1); if (stop || ++k >= n_act) { if (!stop) {
#pragma unroll
          for (int a = 0; a < W; a++) y[a] = fmaf(B0[a], dp, y[a]); } break; }
      }
    } else if (V == 2 && it == 0) {
      // y holds y_after(j-2) when step j starts; row j-1 (buffer `prev`, impulse change dp) is
      // applied to y in the shadow of step j's chain, then row j+1's dot is taken.  Four register
      // buffers in rotation (prev, cur, nxt, loading), the sequence of rows is cyclic.
      float B0[16], B1[16], B2[16], B3[16], l0, l1, l2, l3, c0, c1, c2, c3, g0, g1, g2, g3;
      float dp = 0.f;
#pragma unroll
      for (int a = 0; a < 16; a++) B3[a] = 0.f;           // "row -1"
      l3 = c3 = g3 = 0.f;
      load16(&ent[0][0], B0); l0 = ent[0][16]; c0 = ent[0][17]; g0 = gap4(B0, y, l0);
      load16(&ent[1][0], B1); l1 = ent[1][16]; c1 = ent[1][17]; g1 = 0.f;
      g2 = 0.f; c2 = 0.f; l2 = 0.f;
      auto step = [&](const int j, const float (&prev)[16], const float (&cur)[16], const float lc, const float cc,
                      const float gc, const float (&nxt)[16], const float ln, float &gn, float (&ld)[16], float &ll,
                      float &cl) {
        const int k = j & (NE - 1);
        const float sum = fmaxf(fmaf(-cc, dp, gc), 0.f), d = sum - lc;        // the chain
        ent[k][16] = sum;
        load16(&ent[(j + 2) & (NE - 1)][0], ld); ll = ent[(j + 2) & (NE - 1)][16]; cl = ent[(j + 2) & (NE - 1)][17];
#pragma unroll
        for (int a = 0; a < W; a++) y[a] = fmaf(prev[a], dp, y[a]);           // row j-1 into y
        gn = gap4(nxt, y, ln);                                                 // row j+1 against y_after(j-1)
        const float rr = d * cur[W + 1]; res = fmaxf(res, rr * rr); budget -= fabsf(d);
        dp = d;
      };
      for (int j = 0; j < ITERS * NE; j += 4) {
        step(j, B3, B0, l0, c0, g0, B1, l1, g1, B2, l2, c2);
        step(j + 1, B0, B1, l1, c1, g1, B2, l2, g2, B3, l3, c3);
        step(j + 2, B1, B2, l2, c2, g2, B3, l3, g3, B0, l0, c0);
        step(j + 3, B2, B3, l3, c3, g3, B0, l0, g0, B1, l1, c1);
      }
#pragma unroll
      for (int a = 0; a < W; a++) y[a] = fmaf(B3[a], dp, y[a]);
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) {
    cycles[blockIdx.x] = t1 - t0;
    float s = res + budget;
#pragma unroll
    for (int a = 0; a < W; a++) s += y[a];
    out[blockIdx.x] = s;
  }
}

int main() {
  float h[(NE + 1) * ES];
  srand(1);
  for (int k = 0; k <= NE; k++) {
    float n2 = 0.f;
    for (int a = 0; a < ES; a++) h[k * ES + a] = 0.f;
    for (int a = 0; a < W; a++) { h[k * ES + a] = (float)rand() / RAND_MAX - 0.5f; n2 += h[k * ES + a] * h[k * ES + a]; }
    for (int a = 0; a < W; a++) h[k * ES + a] /= sqrtf(n2);
    h[k * ES + W] = 0.01f * ((float)rand() / RAND_MAX);   // b
    h[k * ES + W + 1] = 1.f;                                 // s
    { unsigned one = 5u; memcpy(&h[k * ES + 18], &one, 4); }   // a non-empty range
  }
  for (int k = 0; k < NE; k++) {                             // c_k = u_k . u_(k-1), cyclic
    float c = 0.f;
    for (int a = 0; a < W; a++) c += h[k * ES + a] * h[((k + NE - 1) % NE) * ES + a];
    h[k * ES + 17] = c;
  }
  float *init, *out; long long *cyc;
  cudaMalloc(&init, sizeof(h)); cudaMalloc(&out, 4096 * 4); cudaMalloc(&cyc, 4096 * 8);
  cudaMemcpy(init, h, sizeof(h), cudaMemcpyHostToDevice);
  const int grids[2] = {1, 148 * 12};
  for (int g = 0; g < 2; g++) {
    for (int v = 0; v < 10; v++) {
      for (int rep = 0; rep < 2; rep++) {
        if (v == 0) rows<0><<<grids[g], 32>>>(init, out, cyc, NE);
        if (v == 1) rows<1><<<grids[g], 32>>>(init, out, cyc, NE);
        if (v == 2) rows<2><<<grids[g], 32>>>(init, out, cyc, NE);
        if (v == 3) rows<3><<<grids[g], 32>>>(init, out, cyc, NE);
        if (v == 4) rows<4><<<grids[g], 32>>>(init, out, cyc, NE);
        if (v == 5) rows<5><<<grids[g], 32>>>(init, out, cyc, NE);
        if (v == 6) rows<6><<<grids[g], 32>>>(init, out, cyc, NE);
        if (v == 7) rows<7><<<grids[g], 32>>>(init, out, cyc, NE);
        if (v == 8) rows<8><<<grids[g], 32>>>(init, out, cyc, NE);
        if (v == 9) rows<9><<<grids[g], 32>>>(init, out, cyc, NE);
        cudaDeviceSynchronize();
      }
      long long c[4096]; float o[4096];
      cudaMemcpy(c, cyc, grids[g] * 8, cudaMemcpyDeviceToHost);
      cudaMemcpy(o, out, grids[g] * 4, cudaMemcpyDeviceToHost);
      double mean = 0; for (int i = 0; i < grids[g]; i++) mean += c[i];
      mean /= grids[g];
      printf("grid %5d variant %d: %.1f cycles per row update (checksum %.6f) %s\n", grids[g], v, mean / (ITERS * NE), o[0],
             cudaGetLastError() == cudaSuccess ? "" : "CUDA ERROR");
    }
  }
  return 0;
}
