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
//                  flat Lambert shade, back-face cull; one 64-byte TriRec per
//                  triangle (affine barycentric and depth equations) plus the robot's screen bounding box per environment.
//   render_raster_kernel   one block per (128 x 8 pixel tile, environment), four
//                  pixels per thread 32 apart in x: the block culls the environment's
//                  triangles against its tile 256 at a time (ballot compaction into
//                  shared memory), every thread depth-tests its pixels against the
//                  survivors, then against the analytic ground plane z = 0
//                  (checkerboard); every store of a warp is one 128-byte piece of a row.
//
// HBM traffic per frame: the 4 W H output bytes; the TriRec table (64 B x
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
  int32_t tiles_per_block;   // vertically adjacent tiles walked by one raster block
};

// One projected triangle.  Barycentric coordinates and depth are affine in the
// pixel position (x, y): lambda_i = e[i].x * x + e[i].y * y + e[i].z, ndc depth
// z = e[0].w * x + e[1].w * y + e[2].w.
struct TriRec {
  float4 e[3];
  int32_t box_lo, box_hi;   // pixel bounding box: (x0 | y0 << 16), (x1 | y1 << 16), int16 each; x1, y1 exclusive
  uint32_t rgba;
  int32_t valid;            // 0: culled (behind the camera, back face, degenerate)
};
static_assert(sizeof(TriRec) == 64, "TriRec is four float4");

__device__ __forceinline__ int32_t pack_box(float x, float y) {
  const int xi = (int)fminf(fmaxf(x, -32768.f), 32767.f), yi = (int)fminf(fmaxf(y, -32768.f), 32767.f);
  return (xi & 0xffff) | (yi << 16);
}

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
  rec.e[0] = rec.e[1] = rec.e[2] = make_float4(0.f, 0.f, 0.f, 0.f);
  rec.box_lo = rec.box_hi = 0;
  rec.rgba = 0;
  rec.valid = 0;
  if (ok) {
    V3<float> nrm = cross(w[1] - w[0], w[2] - w[0]);
    nrm = nrm * m_rsqrt(dot(nrm, nrm) + 1e-30f);
    const float li = R.ambient + R.diffuse * fmaxf(0.f, nrm.x * R.light[0] + nrm.y * R.light[1] + nrm.z * R.light[2]);
    rec.rgba = pack_rgba(tri_rgba[tri * 4 + 0] * li, tri_rgba[tri * 4 + 1] * li, tri_rgba[tri * 4 + 2] * li);
    rec.valid = 1;
    // v1 <-> v2: positive area for the rasteriser; edge i is opposite vertex i
    const float X[3] = {sx[0], sx[2], sx[1]}, Y[3] = {sy[0], sy[2], sy[1]}, Z[3] = {sz[0], sz[2], sz[1]};
    const float inv = -1.f / area;
    float za = 0.f, zb = 0.f, zc = 0.f;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const int a = (i + 1) % 3, b = (i + 2) % 3;
      const float A = -(Y[b] - Y[a]) * inv, B = (X[b] - X[a]) * inv;
      const float C = -(A * X[a] + B * Y[a]);
      rec.e[i].x = A; rec.e[i].y = B; rec.e[i].z = C;
      za += A * Z[i]; zb += B * Z[i]; zc += C * Z[i];
    }
    rec.e[0].w = za; rec.e[1].w = zb; rec.e[2].w = zc;
    const float x0 = fminf(sx[0], fminf(sx[1], sx[2])), x1 = fmaxf(sx[0], fmaxf(sx[1], sx[2]));
    const float y0 = fminf(sy[0], fminf(sy[1], sy[2])), y1 = fmaxf(sy[0], fmaxf(sy[1], sy[2]));
    rec.box_lo = pack_box(floorf(x0), floorf(y0));
    rec.box_hi = pack_box(ceilf(x1), ceilf(y1));
    int32_t *bb = bbox + 4 * slot;
    const float lim = 1e6f;
    atomicMin(bb + 0, (int)floorf(fmaxf(x0, -lim)));
    atomicMin(bb + 1, (int)floorf(fmaxf(y0, -lim)));
    atomicMin(bb + 2, -(int)ceilf(fminf(x1, lim)));
    atomicMin(bb + 3, -(int)ceilf(fminf(y1, lim)));
  }
  recs[(int64_t)slot * R.n_tri + tri] = rec;
}

// kPix pixels per thread, 32 apart in x: every store instruction of a warp is one contiguous
// 128-byte piece of an image row, the per-thread and per-tile bookkeeping is paid once for four
// pixels, and the four pixels' arithmetic is independent work for the scheduler.
// (One pixel per thread = 32-pixel-wide tiles for a handful of frames, which need every block
// they can get; four for batches.)
// AUX: also write getCameraImage's depth buffer and segmentation mask (render() itself discards
// them, like the reference).
template <int kPix, bool AUX>
__global__ void __launch_bounds__(kRasterThreads)
render_raster_kernel(const __grid_constant__ RenderParams R, const TriRec *__restrict__ recs,
                     const int32_t *__restrict__ bbox, uint32_t *__restrict__ out, float *__restrict__ depth_out,
                     int32_t *__restrict__ seg_out) {
  __shared__ float4 se[3][kRasterThreads];
  __shared__ uint32_t s_rgba[kRasterThreads];
  __shared__ int4 s_box[kRasterThreads];   // pixel bounding box of the survivor: x0, x1, y0, y1 (exclusive ends)
  __shared__ int warp_cnt[kRasterThreads / 32];
  const int slot = blockIdx.z;
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  const int tx0 = blockIdx.x * (kTileW * kPix), tx1 = tx0 + kTileW * kPix;
  const int4 bb = *reinterpret_cast<const int4 *>(bbox + 4 * slot);
  const TriRec *mine = recs + (int64_t)slot * R.n_tri;
  const int4 *meta0 = reinterpret_cast<const int4 *>(&mine[t < R.n_tri ? t : 0].box_lo);
  const bool x_touched = bb.x < tx1 && -bb.z > tx0;

  // per-pixel constants of the thread's four columns: pixel centre, and the part of the view ray
  // that depends on x only (ray = ray_z + nx ray_x + ny ray_y)
  float fx[kPix], ax[kPix], ay[kPix], az[kPix];
#pragma unroll
  for (int k = 0; k < kPix; k++) {
    fx[k] = (float)(tx0 + lane + kTileW * k) + 0.5f;
    const float nx = fx[k] * R.ndc_sx - 1.f;
    ax[k] = R.ray_z[0] + nx * R.ray_x[0];
    ay[k] = R.ray_z[1] + nx * R.ray_x[1];
    az[k] = R.ray_z[2] + nx * R.ray_x[2];
  }
  // The checkerboard in cell units.  floor() by the magic-number add: for |c| < 2^21 the float
  // (c - 0.5) + 1.5 * 2^23 is an integer with floor(c) in the low bits of its mantissa
  // (conversions run at an eighth of the FMA rate, and there would be two per pixel); only the
  // parity of the two floors is needed.
  constexpr float kMagic = 12582912.f;
  const float ex = R.eye[0] * R.inv_checker - 0.5f, ey = R.eye[1] * R.inv_checker - 0.5f;

  // several vertically adjacent tiles per block when there are many environments: the short
  // sky / plane tiles amortise the block launch
  for (int sub = 0; sub < R.tiles_per_block; sub++) {
    const int tile_y = blockIdx.y * R.tiles_per_block + sub;
    const int ty0 = tile_y * kTileH, ty1 = ty0 + kTileH;
    if (ty0 >= R.height) break;
    const int py = ty0 + wid;
    const float fy = (float)py + 0.5f;

    float best[kPix];
    uint32_t colour[kPix];
#pragma unroll
    for (int k = 0; k < kPix; k++) { best[k] = 1.f; colour[k] = R.sky_rgba; }   // ndc depth of the far plane

    const bool touched = x_touched && bb.y < ty1 && -bb.w > ty0;
    if (touched) {
      int4 m_next = *meta0;
      for (int base = 0; base < R.n_tri; base += kRasterThreads) {
        const int idx = base + t;
        const int4 m = m_next;
        if (idx + kRasterThreads < R.n_tri)     // next chunk's bounding boxes fly during this chunk's pixels
          m_next = *reinterpret_cast<const int4 *>(&mine[idx + kRasterThreads].box_lo);
        const int x0 = (int)(short)(m.x & 0xffff), y0 = m.x >> 16, x1 = (int)(short)(m.y & 0xffff), y1 = m.y >> 16;
        const bool keep = idx < R.n_tri && m.w != 0 && x0 < tx1 && x1 > tx0 && y0 < ty1 && y1 > ty0;
        const unsigned mask = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_cnt[wid] = __popc(mask);
        __syncthreads();
        int off = 0, total = 0;
#pragma unroll
        for (int k = 0; k < kRasterThreads / 32; k++) {
          const int c = warp_cnt[k];
          off += k < wid ? c : 0;
          total += c;
        }
        if (keep) {
          const int p = off + __popc(mask & ((1u << lane) - 1u));   // triangle order is kept
          const float4 *e = mine[idx].e;
          se[0][p] = e[0]; se[1][p] = e[1]; se[2][p] = e[2];
          s_rgba[p] = (uint32_t)m.z;
          s_box[p] = make_int4(x0, x1, y0, y1);
        }
        __syncthreads();
        for (int k = 0; k < total; k++) {
          // a warp is one image row and four 32-pixel segments of it: a survivor of the tile whose
          // box misses the row is skipped for the price of one shared load, a segment it misses
          // for a compare (all warp-uniform)
          const int4 bx = s_box[k];
          if (py < bx.z || py >= bx.w) continue;
          const float4 e0 = se[0][k], e1 = se[1][k], e2 = se[2][k];
          const uint32_t c = s_rgba[k];
          // the row part of the four affine forms once, the column part per pixel
          const float r0 = fmaf(e0.y, fy, e0.z), r1 = fmaf(e1.y, fy, e1.z), r2 = fmaf(e2.y, fy, e2.z);
          const float rz = fmaf(e1.w, fy, e2.w);
#pragma unroll
          for (int q = 0; q < kPix; q++) {
            const int sx0 = tx0 + kTileW * q;
            if (bx.x < sx0 + kTileW && bx.y > sx0) {
              const float l0 = fmaf(e0.x, fx[q], r0), l1 = fmaf(e1.x, fx[q], r1), l2 = fmaf(e2.x, fx[q], r2);
              if (fminf(l0, fminf(l1, l2)) >= 0.f) {
                const float z = fmaf(e0.w, fx[q], rz);
                if (z < best[q] && z >= -1.f) { best[q] = z; colour[q] = c; }
              }
            }
          }
        }
        __syncthreads();
      }
    }

    // ground plane z = 0 (plane.urdf, solo8_base_env.py:47): analytic ray hit, checkerboard
    const float ny = 1.f - fy * R.ndc_sy;
    const float by = ny * R.ray_y[0], cy = ny * R.ray_y[1], dy = ny * R.ray_y[2];
    const bool row_ok = py < R.height;
    const int64_t row = ((int64_t)slot * R.height + py) * R.width + tx0 + lane;
    uint32_t *out_row = out + row;
#pragma unroll
    for (int k = 0; k < kPix; k++) {
      const bool on_robot = best[k] < 1.f;   // a triangle won so far
      bool on_plane = false;
      const float rz = az[k] + dy;
      const float tt = __fdividef(-R.eye[2], rz);
      const float z = -R.depth_a + R.plane_k * rz;
      if (rz != 0.f && tt > R.near_t && tt < R.far_t && z < best[k]) {
        best[k] = z;
        on_plane = true;
        const float tc = fminf(tt * R.inv_checker, 2.0e6f);      // (cells beyond 2 million: keep the magic add exact)
        const float ca = fmaf(tc, ax[k] + by, ex) + kMagic, cb = fmaf(tc, ay[k] + cy, ey) + kMagic;
        colour[k] = ((__float_as_int(ca) ^ __float_as_int(cb)) & 1) ? R.plane_rgba_b : R.plane_rgba_a;
      }
      if (row_ok && tx0 + lane + kTileW * k < R.width) {
        out_row[kTileW * k] = colour[k];
        if (AUX) {
          // getCameraImage's other two images: OpenGL depth in [0, 1] (1 = far plane / nothing hit)
          // and the body id (plane 0, robot 1, nothing -1)
          if (depth_out) depth_out[row + kTileW * k] = 0.5f * best[k] + 0.5f;
          if (seg_out) seg_out[row + kTileW * k] = on_plane ? 0 : (on_robot ? 1 : -1);
        }
      }
    }
  }
}

}  // namespace solo
#endif  // SOLO_RENDER_CUH_
