// solo_render.cuh -- headless rasteriser behind Solo8BaseEnv.render()
// (gym_solo/envs/solo8_base_env.py:116-142: computeProjectionMatrixFOV,
// computeViewMatrixFromYawPitchRoll, getCameraImage -> (h, w, 4) RGBA).
//
// No graphics API is involved: the image is a device buffer written by two
// launches on the caller's stream.
//
//   render_project_kernel  one thread per (environment, triangle): body poses
//                  by forward kinematics from the SoA state (thirteen threads of
//                  the block, shared memory), vertices to pixel coordinates,
//                  flat Lambert shade, back-face cull; one 48-byte TriRec per
//                  triangle plus the robot's screen bounding box per environment.
//   render_raster_kernel   one block per (32 x 8 pixel tile, environment): the
//                  block culls the environment's triangles against its tile 256
//                  at a time (ballot compaction into shared memory), every thread
//                  depth-tests its pixel against the survivors, then against the
//                  analytic ground plane z = 0 (checkerboard); one coalesced
//                  128-byte row of RGBA per warp.
//
// HBM traffic per frame: the 4 W H output bytes; the TriRec table (48 B x
// n_tri per environment) is written once and re-read from L2 by the tiles inside
// the robot's bounding box only.
#ifndef SOLO_RENDER_CUH_
#define SOLO_RENDER_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include "solo_core.h"

namespace solo {

constexpr int kTileW = 32, kTileH = 8, kRasterThreads = kTileW * kTileH;
constexpr int kProjectThreads = 128;

struct RenderParams {
  float vp[16];            // proj * view, row-major: clip = vp * (x y z 1)
  float eye[3];            // camera position, world
  float ray_x[3];          // camera right / proj xScale
  float ray_y[3];          // camera up / proj yScale
  float ray_z[3];          // camera forward
  float depth_a;           // ground plane: ndc z = -depth_a + plane_k * ray.z
  float plane_k;           //   (= -proj[2][3] / eye.z; t = -eye.z / ray.z is the distance along ray_z)
  float near_t, far_t;
  float ndc_sx, ndc_sy;    // 2 / width, 2 / height
  float light[3];          // unit vector towards the light
  float ambient, diffuse;
  uint32_t sky_rgba, plane_rgba_a, plane_rgba_b;   // packed, the plane colours already shaded
  float inv_checker;
  int32_t width, height, n_tri, env_begin, env_count;
};

struct TriRec {
  float4 a;   // x0 y0 x1 y1   (pixel coordinates, y down)
  float4 b;   // x2 y2 z0 z1   (z = ndc depth)
  float4 c;   // z2, 1 / area (0: culled), rgba8 bits, unused
};

__host__ __device__ __forceinline__ uint32_t pack_rgba(float r, float g, float b) {
  const uint32_t R = (uint32_t)(fminf(fmaxf(r, 0.f), 1.f) * 255.f + 0.5f);
  const uint32_t G = (uint32_t)(fminf(fmaxf(g, 0.f), 1.f) * 255.f + 0.5f);
  const uint32_t B = (uint32_t)(fminf(fmaxf(b, 0.f), 1.f) * 255.f + 0.5f);
  return R | (G << 8) | (B << 16) | 0xff000000u;
}

// bbox[env] = (min x, min y, -max x, -max y) in pixels, all reduced with atomicMin
__global__ void __launch_bounds__(kProjectThreads)
render_project_kernel(const __grid_constant__ ModelT<float> M, const __grid_constant__ RenderParams R,
                      const float *__restrict__ state, const int32_t n_envs,
                      const int32_t *__restrict__ tri_body, const float *__restrict__ tri_verts,
                      const float *__restrict__ tri_rgba, TriRec *__restrict__ recs, int32_t *__restrict__ bbox) {
  __shared__ float pose[kMaxBodies][12];   // rotation (row-major 9) + origin 3, world
  const int slot = blockIdx.y;
  const int64_t env = R.env_begin + slot;
  const int t = threadIdx.x;
  if (t <= M.ndof) {
    float q4[4];
    for (int i = 0; i < 4; i++) q4[i] = state[(3 + i) * (int64_t)n_envs + env];
    M3<float> Rw = quat_to_mat(q4);
    V3<float> o = mk(state[env], state[(int64_t)n_envs + env], state[2 * (int64_t)n_envs + env]);
    if (t > 0) {
      const int d_target = t - 1;
      int leg = 0;
      for (int l = 0; l < M.n_legs; l++)
        if (d_target >= M.dof0[l] && d_target < M.dof0[l] + M.chain_len[l]) leg = l;
      for (int d = M.dof0[leg]; d <= d_target; d++) {
        M3<float> Jr;
        for (int k = 0; k < 9; k++) Jr.m[k] = M.jrot[d][k];
        o = o + mul(Rw, ld3(M.jorigin[d]));
        const float q = state[(kStateBase + d) * (int64_t)n_envs + env];
        Rw = mul(mul(Rw, Jr), axis_angle(ld3(M.jaxis[d]), q));
      }
    }
    for (int k = 0; k < 9; k++) pose[t][k] = Rw.m[k];
    pose[t][9] = o.x; pose[t][10] = o.y; pose[t][11] = o.z;
  }
  __syncthreads();
  const int tri = blockIdx.x * kProjectThreads + t;
  if (tri >= R.n_tri) return;
  const int b = tri_body[tri];
  M3<float> Rb;
  for (int k = 0; k < 9; k++) Rb.m[k] = pose[b][k];
  const V3<float> ob = mk(pose[b][9], pose[b][10], pose[b][11]);
  V3<float> w[3];
  float sx[3], sy[3], sz[3];
  bool ok = true;
  for (int v = 0; v < 3; v++) {
    w[v] = ob + mul(Rb, ld3(tri_verts + (int64_t)tri * 9 + 3 * v));
    const float cx = R.vp[0] * w[v].x + R.vp[1] * w[v].y + R.vp[2] * w[v].z + R.vp[3];
    const float cy = R.vp[4] * w[v].x + R.vp[5] * w[v].y + R.vp[6] * w[v].z + R.vp[7];
    const float cz = R.vp[8] * w[v].x + R.vp[9] * w[v].y + R.vp[10] * w[v].z + R.vp[11];
    const float cw = R.vp[12] * w[v].x + R.vp[13] * w[v].y + R.vp[14] * w[v].z + R.vp[15];
    ok = ok && cw > 1e-6f;
    const float iw = 1.f / cw;
    sx[v] = (cx * iw * 0.5f + 0.5f) * (float)R.width;
    sy[v] = (0.5f - cy * iw * 0.5f) * (float)R.height;
    sz[v] = cz * iw;
  }
  // a front face is counter-clockwise seen from outside: negative area with y pointing down
  float area = (sx[1] - sx[0]) * (sy[2] - sy[0]) - (sx[2] - sx[0]) * (sy[1] - sy[0]);
  ok = ok && area < -1e-12f;
  TriRec rec;
  rec.a = make_float4(sx[0], sy[0], sx[2], sy[2]);      // v1 <-> v2: positive area for the rasteriser
  rec.b = make_float4(sx[1], sy[1], sz[0], sz[2]);
  rec.c = make_float4(sz[1], 0.f, 0.f, 0.f);
  if (ok) {
    V3<float> nrm = cross(w[1] - w[0], w[2] - w[0]);
    nrm = nrm * m_rsqrt(dot(nrm, nrm) + 1e-30f);
    const float li = R.ambient + R.diffuse * fmaxf(0.f, nrm.x * R.light[0] + nrm.y * R.light[1] + nrm.z * R.light[2]);
    rec.c.y = -1.f / area;
    rec.c.z = __uint_as_float(pack_rgba(tri_rgba[tri * 4 + 0] * li, tri_rgba[tri * 4 + 1] * li,
                                        tri_rgba[tri * 4 + 2] * li));
    int32_t *bb = bbox + 4 * slot;
    const float x0 = fminf(sx[0], fminf(sx[1], sx[2])), x1 = fmaxf(sx[0], fmaxf(sx[1], sx[2]));
    const float y0 = fminf(sy[0], fminf(sy[1], sy[2])), y1 = fmaxf(sy[0], fmaxf(sy[1], sy[2]));
    const float lim = 1e6f;
    atomicMin(bb + 0, (int)floorf(fmaxf(x0, -lim)));
    atomicMin(bb + 1, (int)floorf(fmaxf(y0, -lim)));
    atomicMin(bb + 2, -(int)ceilf(fminf(x1, lim)));
    atomicMin(bb + 3, -(int)ceilf(fminf(y1, lim)));
  }
  recs[(int64_t)slot * R.n_tri + tri] = rec;
}

__global__ void __launch_bounds__(kRasterThreads)
render_raster_kernel(const __grid_constant__ RenderParams R, const TriRec *__restrict__ recs,
                     const int32_t *__restrict__ bbox, uint32_t *__restrict__ out) {
  __shared__ float4 sa[kRasterThreads], sb[kRasterThreads], sc[kRasterThreads];
  __shared__ int warp_cnt[kRasterThreads / 32];
  const int slot = blockIdx.z;
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  const int tile_x = blockIdx.x, tile_y = blockIdx.y;
  const int px = tile_x * kTileW + lane, py = tile_y * kTileH + wid;
  const float fx = (float)px + 0.5f, fy = (float)py + 0.5f;
  const float tx0 = (float)(tile_x * kTileW), tx1 = tx0 + (float)kTileW;
  const float ty0 = (float)(tile_y * kTileH), ty1 = ty0 + (float)kTileH;

  float best = 1.f;                       // ndc depth of the far plane
  uint32_t colour = R.sky_rgba;

  const int4 bb = *reinterpret_cast<const int4 *>(bbox + 4 * slot);
  const bool touched = (float)bb.x < tx1 && (float)(-bb.z) > tx0 && (float)bb.y < ty1 && (float)(-bb.w) > ty0;
  if (touched) {
    const TriRec *mine = recs + (int64_t)slot * R.n_tri;
    for (int base = 0; base < R.n_tri; base += kRasterThreads) {
      const int idx = base + t;
      TriRec rec;
      bool keep = false;
      if (idx < R.n_tri) {
        rec = mine[idx];
        if (rec.c.y != 0.f) {
          const float x0 = fminf(rec.a.x, fminf(rec.a.z, rec.b.x)), x1 = fmaxf(rec.a.x, fmaxf(rec.a.z, rec.b.x));
          const float y0 = fminf(rec.a.y, fminf(rec.a.w, rec.b.y)), y1 = fmaxf(rec.a.y, fmaxf(rec.a.w, rec.b.y));
          keep = x0 < tx1 && x1 > tx0 && y0 < ty1 && y1 > ty0;
        }
      }
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      if (lane == 0) warp_cnt[wid] = __popc(m);
      __syncthreads();
      int off = 0, total = 0;
#pragma unroll
      for (int k = 0; k < kRasterThreads / 32; k++) {
        const int c = warp_cnt[k];
        off += k < wid ? c : 0;
        total += c;
      }
      if (keep) {
        const int p = off + __popc(m & ((1u << lane) - 1u));   // triangle order is kept
        sa[p] = rec.a; sb[p] = rec.b; sc[p] = rec.c;
      }
      __syncthreads();
      for (int k = 0; k < total; k++) {
        const float4 a = sa[k], b = sb[k], c = sc[k];
        // edge functions of (v0, v1, v2) = (a.xy, a.zw, b.xy), positive inside
        const float w0 = (b.x - a.z) * (fy - a.w) - (b.y - a.w) * (fx - a.z);
        const float w1 = (a.x - b.x) * (fy - b.y) - (a.y - b.y) * (fx - b.x);
        const float w2 = (a.z - a.x) * (fy - a.y) - (a.w - a.y) * (fx - a.x);
        if (w0 >= 0.f && w1 >= 0.f && w2 >= 0.f) {
          const float z = (w0 * b.z + w1 * b.w + w2 * c.x) * c.y;
          if (z < best && z >= -1.f) { best = z; colour = __float_as_uint(c.z); }
        }
      }
      __syncthreads();
    }
  }

  // ground plane z = 0 (plane.urdf, solo8_base_env.py:47): analytic ray hit, checkerboard
  {
    const float nx = fx * R.ndc_sx - 1.f, ny = 1.f - fy * R.ndc_sy;
    const float rz = R.ray_z[2] + nx * R.ray_x[2] + ny * R.ray_y[2];
    const float tt = -R.eye[2] / rz;
    const float z = -R.depth_a + R.plane_k * rz;
    if (rz != 0.f && tt > R.near_t && tt < R.far_t && z < best) {
      const float hx = R.eye[0] + tt * (R.ray_z[0] + nx * R.ray_x[0] + ny * R.ray_y[0]);
      const float hy = R.eye[1] + tt * (R.ray_z[1] + nx * R.ray_x[1] + ny * R.ray_y[1]);
      const int cell = (int)floorf(hx * R.inv_checker) + (int)floorf(hy * R.inv_checker);
      colour = (cell & 1) ? R.plane_rgba_b : R.plane_rgba_a;
    }
  }
  if (px < R.width && py < R.height)
    out[((int64_t)slot * R.height + py) * R.width + px] = colour;
}

}  // namespace solo
#endif  // SOLO_RENDER_CUH_
