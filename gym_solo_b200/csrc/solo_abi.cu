// solo_abi.cu -- the C-ABI of include/solo8_b200.h over the sm_100a kernels.
// No torch types, no CPU fallback: without a CUDA device solo_create fails with
// SOLO_ENOGPU and nothing else can be called.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/solo8_b200.h"
#include "solo_convert.h"
#include "solo_kernels.cuh"
#include "solo_render.cuh"
#include "solo_setup_legpar.cuh"
#include "solo_solve_warp.cuh"

using namespace solo;

struct SoloHandle_ {
  KParams K;
  Buffers B;
  int device;
  int ch;                 // kernel instantiation: max chain length (2 or 3)
  int state_dim;
  int obs_dim;
  bool legpar;            // setup phase with one lane per (environment, leg): latency-bound batch sizes
  int block;              // threads per block of the one-thread-per-env kernels
  int persist_grid;       // grid of the persistent solver build: one resident wave of one-warp blocks
  int solve_mode;         // solver build: 0 (and 1, kept as an alias) = one block per env; 2 = persistent warps
                          // (SOLO_OPT_SOLVER_OCCUPANCY)
  bool warm_finish;       // finish_kernel runs a dry pass behind the solver (small batches: fetch bound)
  uint32_t motor_mask;    // live motor rows (0 in torque mode)
  bool program_set;
  bool bound;
  // staging (solo_step_host)
  float *st_actions, *st_obs, *st_reward;
  uint8_t *st_done;
  int64_t launches;
  // one env.step as ONE graph launch: the (setup, solve, finish) x substeps chain of solo_step is
  // captured once per set of output buffers and replayed; only the action pointer of the setup
  // nodes is patched per call.  Anything baked into the kernel parameters (bindings, program)
  // bumps `epoch` and retires the captured graphs.
  bool use_graph;
  uint64_t epoch;
  std::vector<struct StepGraph *> graphs;
  cudaStream_t gstream;   // capture / replay stream when the caller's is the legacy default stream
  // zero-copy host entry: device aliases of the caller's pinned result buffers (solo_step_host)
  bool zero_copy;
  const void *map_key[3];
  void *map_dev[3];
  // render(): triangle mesh in body coordinates (device, caller-owned)
  const int32_t *tri_body;
  const float *tri_verts, *tri_rgba;
  int32_t n_tri;
  // optional live per-phase timing (bench.py roofline): 4 events per physics step
  bool timing;
  std::vector<cudaEvent_t> ev;
  size_t ev_used;
  double phase_ms[3];
  int64_t phase_steps;
  std::string err;
};

static thread_local std::string g_err;

static int fail(SoloHandle h, int code, const std::string &msg) {
  if (h) h->err = msg;
  g_err = msg;
  return code;
}
static int cuda_fail(SoloHandle h, cudaError_t e, const char *what) {
  return fail(h, SOLO_ECUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

struct DeviceGuard {
  int prev;
  bool ok;
  explicit DeviceGuard(int dev) : prev(-1), ok(true) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    int cur = -1;
    if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
  }
};

static int pick_block(int n_envs) {
#ifdef SOLO_EXPERIMENTS
  if (const char *e = getenv("SOLO_BLOCK")) {   // A/B builds only (tools/): never in the product library
    int b = atoi(e);
    if (b == 32 || b == 64 || b == 128) return b;
  }
#endif
  // 32 environments per block = 128 threads in the leg-parallel setup kernel (158 registers: three
  // blocks per SM instead of one 256-thread block) and one warp per block in the finish kernel.
  // Measured against 64: setup 0.141 -> 0.119 ms at 65536 envs, 0.077 -> 0.069 at 32768, +3 % at
  // 16384, the same below; finish unchanged.
  (void)n_envs;
  return 32;
}

constexpr int kHiOccEnvs = 10240;   // below this finish_kernel's dry warm-up pass pays (SOLO_OPT_FINISH_WARMUP)
#ifndef SOLO_PERSIST_ENVS
#define SOLO_PERSIST_ENVS 28672      // from here the persistent solver build (measured: +2 % at 24576, -3 % at 32768, -10 % at 65536)
#endif
// Register cap of the one-block-per-environment build: chains of 3 joints (CH = 3) are held to 168
// registers (10 blocks per SM; measured against uncapped and against 128 registers, profiles/README.md);
// CH = 2 needs 127 and is limited by shared memory.  The persistent build stays uncapped.
#define SOLO_MINB0(CH) ((CH) == 3 ? 10 : 0)
static int auto_solve_mode(int n) { return n >= SOLO_PERSIST_ENVS ? 2 : 0; }

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct ScratchLayout {
  size_t cand, yh, fac, kin, work, total;
};
static ScratchLayout scratch_layout(const SoloHandle h) {
  const size_t n = (size_t)h->B.n;
  const size_t facs = h->ch == 2 ? FactorLayout<2>::kSize : FactorLayout<3>::kSize;
  const size_t kins = h->ch == 2 ? KinLayout<2>::kSize : KinLayout<3>::kSize;
  ScratchLayout L;
  size_t off = 0;
  L.cand = off; off = align256(off + sizeof(float) * kCandStride * kMaxCand);
  L.yh = off;   off = align256(off + sizeof(float) * (6 + kLegs * h->ch + 4) * n);
  L.fac = off;  off = align256(off + sizeof(float) * facs * n);
  L.kin = off;  off = align256(off + sizeof(float) * kins * n);
  L.work = off; off = align256(off + 2 * sizeof(int32_t));
  L.total = off;
  return L;
}

static size_t tile_smem(const SoloHandle h) {
  int k = h->K.M.n_api > h->obs_dim ? h->K.M.n_api : h->obs_dim;
  return (size_t)h->block * (size_t)(k | 1) * sizeof(float);
}

// setup_legpar_kernel: the action tile + the block's KinLayout records (bulk-stored at the end)
static size_t setup_smem(const SoloHandle h) {
  const size_t tile = (((size_t)h->block * (size_t)(h->K.M.n_api | 1) + 3) & ~(size_t)3) * sizeof(float);
  const size_t rec = (size_t)h->block * (h->ch == 2 ? KinLayout<2>::kSize : KinLayout<3>::kSize) * sizeof(float);
  return tile + rec;
}

template <typename F>
static int opt_in_smem(SoloHandle h, F *kernel, size_t bytes) {
  if (bytes <= 48 * 1024) return SOLO_OK;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaFuncSetAttribute");
  return SOLO_OK;
}

// SoloCamera (column-major view / proj) -> the constants the render kernels read
static std::string make_render_params(const SoloCamera &cam, RenderParams *out) {
  RenderParams &R = *out;
  memset(&R, 0, sizeof R);
  auto V = [&](int r, int c) { return (double)cam.view[c * 4 + r]; };
  auto P = [&](int r, int c) { return (double)cam.proj[c * 4 + r]; };
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) {
      double a = 0;
      for (int k = 0; k < 4; k++) a += P(r, k) * V(k, c);
      R.vp[r * 4 + c] = (float)a;
    }
  const double xs = P(0, 0), ys = P(1, 1), A = P(2, 2), Bz = P(2, 3);
  if (!(xs > 0) || !(ys > 0) || P(3, 2) != -1.0 || !(Bz < 0) || A == 1.0 || A == -1.0)
    return "solo_render: proj is not a perspective projection (computeProjectionMatrixFOV)";
  for (int i = 0; i < 3; i++) {
    // view = [Rv | t]: eye = -Rv^T t; rows of Rv are right, up, -forward
    R.eye[i] = (float)-(V(0, i) * V(0, 3) + V(1, i) * V(1, 3) + V(2, i) * V(2, 3));
    R.ray_x[i] = (float)(V(0, i) / xs);
    R.ray_y[i] = (float)(V(1, i) / ys);
    R.ray_z[i] = (float)-V(2, i);
  }
  R.depth_a = (float)A;
  R.plane_k = R.eye[2] != 0.f ? (float)(-Bz / (double)R.eye[2]) : 0.f;
  R.near_t = (float)(Bz / (A - 1.0));
  R.far_t = (float)(Bz / (A + 1.0));
  R.ndc_sx = (float)(2.0 / cam.width);
  R.ndc_sy = (float)(2.0 / cam.height);
  double l2 = 0;
  for (int i = 0; i < 3; i++) l2 += (double)cam.light_direction[i] * cam.light_direction[i];
  if (!(l2 > 0)) return "solo_render: light_direction is zero";
  for (int i = 0; i < 3; i++) R.light[i] = (float)(cam.light_direction[i] / sqrt(l2));
  R.ambient = cam.ambient;
  R.diffuse = cam.diffuse;
  // the plane's normal is +z: one shade for the whole checkerboard
  const float li = R.ambient + R.diffuse * fmaxf(0.f, R.light[2]);
  R.sky_rgba = pack_rgba(cam.background[0], cam.background[1], cam.background[2]);
  R.plane_rgba_a = pack_rgba(cam.plane_rgb_a[0] * li, cam.plane_rgb_a[1] * li, cam.plane_rgb_a[2] * li);
  R.plane_rgba_b = pack_rgba(cam.plane_rgb_b[0] * li, cam.plane_rgb_b[1] * li, cam.plane_rgb_b[2] * li);
  if (!(cam.checker_size > 0)) return "solo_render: checker_size must be positive";
  R.inv_checker = 1.f / cam.checker_size;
  R.width = cam.width;
  R.height = cam.height;
  return "";
}

// ---- launches ------------------------------------------------------------------------
// pdl: programmatic dependent launch -- the kernel may be scheduled while its predecessor in
// the stream is still running; it calls cudaGridDependencySynchronize() before it touches
// anything the predecessor writes (solve_warp_kernel, finish_kernel).
template <typename... KArgs, typename... Args>
static cudaError_t launch_ex(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t st, bool pdl,
                             Args &&...args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof cfg);
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// One resident wave of the persistent solver build (what the occupancy calculator says fits, times
// the SM count), or one block per environment when that is fewer.
template <int CH>
static int persist_grid(SoloHandle h) {
  if (h->persist_grid == 0) {
    int per_sm = 0, sms = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, solve_warp_kernel<CH, false, 0, true>, 32, 0) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device) != cudaSuccess || per_sm < 1 || sms < 1) {
      cudaGetLastError();
      per_sm = 12; sms = 148;
    }
    h->persist_grid = per_sm * sms;
  }
  return h->B.n < h->persist_grid ? h->B.n : h->persist_grid;
}

// Argument block of one captured setup node (cudaGraphExecKernelNodeSetParams wants every
// argument again when one changes).
struct SetupArgs {
  KParams K;
  Buffers B;
  const float *actions;
  int broadcast;
  float scale;
  void *ptrs[5];
  SetupArgs() { ptrs[0] = &K; ptrs[1] = &B; ptrs[2] = &actions; ptrs[3] = &broadcast; ptrs[4] = &scale; }
  SetupArgs(const SetupArgs &o) : K(o.K), B(o.B), actions(o.actions), broadcast(o.broadcast), scale(o.scale) {
    ptrs[0] = &K; ptrs[1] = &B; ptrs[2] = &actions; ptrs[3] = &broadcast; ptrs[4] = &scale;
  }
};

struct StepGraph {
  cudaGraph_t graph;
  cudaGraphExec_t exec;
  std::vector<cudaGraphNode_t> setup_nodes;
  std::vector<cudaKernelNodeParams> setup_params;
  SetupArgs *args;            // one block shared by the setup nodes of the substeps
  const float *actions;       // what the nodes currently point at
  float scale;
  float *obs, *reward, *terminal;
  uint8_t *done;
  uint64_t epoch, used;
};

static void drop_graphs(SoloHandle h) {
  for (StepGraph *g : h->graphs) {
    if (g->exec) cudaGraphExecDestroy(g->exec);
    if (g->graph) cudaGraphDestroy(g->graph);
    delete g->args;
    delete g;
  }
  h->graphs.clear();
}
static void touch(SoloHandle h) { h->epoch++; }   // kernel parameters changed: captured graphs are stale

extern "C" {

const char *solo_version(void) { return "solo8_b200 0.1 (sm_100a)"; }

const char *solo_last_error(SoloHandle h) { return h ? h->err.c_str() : g_err.c_str(); }

int solo_create(const SoloModel *model, const SoloConfig *config, int32_t n_envs, int32_t device,
                SoloHandle *out) {
  if (!model || !config || !out) return fail(nullptr, SOLO_EINVAL, "null argument");
  *out = nullptr;
  if (n_envs < 1) return fail(nullptr, SOLO_EINVAL, "n_envs must be >= 1");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count < 1) {
    cudaGetLastError();
    return fail(nullptr, SOLO_ENOGPU, "no CUDA device: the Solo 8 step path has no CPU fallback");
  }
  if (device < 0 || device >= count) return fail(nullptr, SOLO_EINVAL, "device index out of range");
  SoloHandle h = new SoloHandle_();
  memset(&h->K, 0, sizeof h->K);
  memset(&h->B, 0, sizeof h->B);
  std::string msg = convert_model<float>(*model, config->plane_friction, &h->K.M);
  if (msg.empty()) msg = convert_config<float>(*config, &h->K.C);
  if (!msg.empty()) { delete h; return fail(nullptr, SOLO_EINVAL, msg); }
  h->device = device;
  h->ch = max_chain(*model) <= 2 ? 2 : 3;
  h->state_dim = SOLO_STATE_DIM(model->ndof);
  h->obs_dim = 0;
  h->block = pick_block(n_envs);
  {
    // setup phase: one lane per (environment, leg) -- faster at every batch size measured
    // (4096: 0.031 -> 0.019 ms, 65536: 0.159 -> 0.138 ms).  The one-thread-per-environment
    // kernel exists only in -DSOLO_EXPERIMENTS builds (SOLO_SETUP=serial, A/B runs).
    h->legpar = true;
#ifdef SOLO_EXPERIMENTS
    if (const char *e = getenv("SOLO_SETUP")) h->legpar = strcmp(e, "serial") != 0;
#endif
    if (h->legpar && h->block > 64) h->block = 64;   // 4 lanes per environment, 256-thread blocks
  }
  // Up to a few waves of solver blocks the launch lasts as long as its slowest environment: one block per
  // environment; with many waves issue slots are the limit and the persistent build wins (32768: -2 %, 65536: -9 %)
  h->solve_mode = auto_solve_mode(n_envs);
  h->persist_grid = 0;
  h->warm_finish = n_envs < kHiOccEnvs;
  h->motor_mask = 0;
  if (config->control_mode == SOLO_CONTROL_POSITION)
    for (int l = 0; l < model->n_legs; l++)
      for (int i = 0; i < model->chain_len[l]; i++) h->motor_mask |= 1u << (l * h->ch + i);
  h->program_set = false;
  h->bound = false;
  h->B.n = n_envs;
  h->st_actions = h->st_obs = h->st_reward = nullptr;
  h->st_done = nullptr;
  h->launches = 0;
  h->use_graph = true;
  h->epoch = 1;
  h->gstream = nullptr;
  h->zero_copy = true;
  h->map_key[0] = h->map_key[1] = h->map_key[2] = nullptr;
  h->map_dev[0] = h->map_dev[1] = h->map_dev[2] = nullptr;
  h->tri_body = nullptr;
  h->tri_verts = h->tri_rgba = nullptr;
  h->n_tri = 0;
  h->timing = false;
  h->ev_used = 0;
  h->phase_ms[0] = h->phase_ms[1] = h->phase_ms[2] = 0;
  h->phase_steps = 0;
  {
    DeviceGuard g(device);
    if (!g.ok) { delete h; return fail(nullptr, SOLO_ECUDA, "cudaSetDevice failed"); }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess && prop.major < 10) {
      delete h;
      return fail(nullptr, SOLO_ENOGPU, "device is not sm_100 class: this library is built for sm_100a only");
    }
    // the setup kernel's shared memory (records of a block) is above 48 KB for 3-joint chains
    int rc = h->ch == 2 ? opt_in_smem(h, setup_legpar_kernel<2>, setup_smem(h))
                        : opt_in_smem(h, setup_legpar_kernel<3>, setup_smem(h));
    if (rc) { std::string m = h->err; delete h; return fail(nullptr, rc, m); }
  }
  *out = h;
  return SOLO_OK;
}

int solo_destroy(SoloHandle h) {
  if (!h) return SOLO_EINVAL;
  {
    DeviceGuard g(h->device);
    drop_graphs(h);
    if (h->gstream) cudaStreamDestroy(h->gstream);
    for (cudaEvent_t e : h->ev) cudaEventDestroy(e);
  }
  delete h;
  return SOLO_OK;
}

int solo_set_program(SoloHandle h, const SoloProgram *program) {
  if (!h || !program) return fail(h, SOLO_EINVAL, "null argument");
  int dim = 0;
  std::string msg = check_program(*program, h->K.M.n_api, &dim);
  if (!msg.empty()) return fail(h, SOLO_EINVAL, msg);
  memcpy(&h->K.P, program, sizeof h->K.P);
  touch(h);
  h->obs_dim = dim;
  h->B.obs_dim = dim;
  h->program_set = true;
  DeviceGuard g(h->device);
  const size_t sm = tile_smem(h);
  int rc = SOLO_OK;
  if (h->ch == 2) {
#ifdef SOLO_EXPERIMENTS
    if (!rc) rc = opt_in_smem(h, setup_kernel<2>, sm);
#endif
    if (!rc) rc = opt_in_smem(h, setup_legpar_kernel<2>, setup_smem(h));
    if (!rc) rc = opt_in_smem(h, finish_kernel<2>, sm);
    if (!rc) rc = opt_in_smem(h, observe_kernel<2>, sm);
  } else {
#ifdef SOLO_EXPERIMENTS
    if (!rc) rc = opt_in_smem(h, setup_kernel<3>, sm);
#endif
    if (!rc) rc = opt_in_smem(h, setup_legpar_kernel<3>, setup_smem(h));
    if (!rc) rc = opt_in_smem(h, finish_kernel<3>, sm);
    if (!rc) rc = opt_in_smem(h, observe_kernel<3>, sm);
  }
  return rc;
}

int solo_obs_dim(SoloHandle h) { return h ? h->obs_dim : SOLO_EINVAL; }

int64_t solo_scratch_bytes(SoloHandle h) {
  if (!h) return SOLO_EINVAL;
  return (int64_t)scratch_layout(h).total;   // per-env records, factors, yh, support-point table
}

int solo_bind(SoloHandle h, float *state_dev, int32_t *counters_dev, float *snapshot_dev, float *stats_dev,
              void *scratch_dev) {
  if (!h || !state_dev || !counters_dev || !snapshot_dev || !scratch_dev) return fail(h, SOLO_EINVAL, "null buffer");
  if (((uintptr_t)scratch_dev & 255) != 0) return fail(h, SOLO_EINVAL, "scratch must be 256-byte aligned");
  {
    const ScratchLayout L = scratch_layout(h);
    char *base = (char *)scratch_dev;
    h->B.cand_tab = (const float *)(base + L.cand);
    h->B.yh = (float *)(base + L.yh);
    h->B.fac = (float *)(base + L.fac);
    h->B.kin = (float *)(base + L.kin);
    h->B.work = (int32_t *)(base + L.work);
    // support-point table of the solver warp: xyz, friction, leg (-1 base), depth in the chain
    std::vector<float> tab_v(kCandStride * kMaxCand, 0.f);   // per call: handles may be bound from several threads
    float *tab = tab_v.data();
    const ModelT<float> &M = h->K.M;
    for (int body = 0; body <= M.ndof; body++) {
      int leg = -1, depth = 0;
      for (int l = 0; l < M.n_legs; l++)
        if (body >= 1 && body - 1 >= M.dof0[l] && body - 1 < M.dof0[l] + M.chain_len[l]) { leg = l; depth = body - 1 - M.dof0[l]; }
      for (int c = M.cand_begin[body]; c < M.cand_begin[body + 1]; c++) {
        for (int k = 0; k < 3; k++) tab[kCandStride * c + k] = M.cand_pos[c][k];
        tab[kCandStride * c + 3] = M.cand_mu[c];
        tab[kCandStride * c + 4] = (float)leg;
        tab[kCandStride * c + 5] = (float)depth;
      }
    }
    DeviceGuard g(h->device);
    cudaError_t e = cudaMemcpy(base + L.cand, tab, sizeof(float) * tab_v.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return cuda_fail(h, e, "upload of the support-point table");
    e = cudaMemset(base + L.work, 0, 2 * sizeof(int32_t));   // the persistent solver's counters rest at zero
    if (e != cudaSuccess) return cuda_fail(h, e, "cudaMemset of the solver's work counters");
  }
  touch(h);
  h->B.state = state_dev;
  h->B.counters = counters_dev;
  h->B.snapshot = snapshot_dev;
  h->B.stats = stats_dev;
  h->bound = true;
  return SOLO_OK;
}

int solo_snapshot_layout(SoloHandle h, int32_t per_env) {
  if (!h) return SOLO_EINVAL;
  touch(h);
  h->B.snapshot_per_env = per_env ? 1 : 0;
  return SOLO_OK;
}

int solo_bind_diag(SoloHandle h, unsigned long long *diag_dev) {
  if (!h) return SOLO_EINVAL;
  touch(h);
  h->B.diag = diag_dev;
  return SOLO_OK;
}

int solo_bind_link_friction(SoloHandle h, const float *link_friction_dev) {
  if (!h) return SOLO_EINVAL;
  touch(h);
  h->B.link_friction = link_friction_dev;
  return SOLO_OK;
}

int solo_bind_base_mass_scale(SoloHandle h, const float *scale_dev) {
  if (!h) return SOLO_EINVAL;
  touch(h);
  h->B.base_mass_scale = scale_dev;
  return SOLO_OK;
}

int solo_bind_debug(SoloHandle h, float *normal_force_dev, float *motor_torque_dev) {
  if (!h) return SOLO_EINVAL;
  touch(h);
  h->B.dbg_normal_force = normal_force_dev;
  h->B.dbg_motor_torque = motor_torque_dev;
  return SOLO_OK;
}

int solo_bind_staging(SoloHandle h, float *actions_dev, float *obs_dev, float *reward_dev, uint8_t *done_dev) {
  if (!h || !actions_dev || !obs_dev || !reward_dev || !done_dev) return fail(h, SOLO_EINVAL, "null buffer");
  h->st_actions = actions_dev; h->st_obs = obs_dev; h->st_reward = reward_dev; h->st_done = done_dev;
  return SOLO_OK;
}

static int check_launch(SoloHandle h, const char *what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(h, e, what);
  h->launches++;
  return SOLO_OK;
}

// Fold the recorded phase events into the running sums (synchronises on them).
static int drain_phase_events(SoloHandle h) {
  for (size_t i = 0; i + 4 <= h->ev_used; i += 4) {
    cudaError_t e = cudaEventSynchronize(h->ev[i + 3]);
    if (e != cudaSuccess) return cuda_fail(h, e, "cudaEventSynchronize");
    for (int p = 0; p < 3; p++) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, h->ev[i + p], h->ev[i + p + 1]);
      h->phase_ms[p] += ms;
    }
    h->phase_steps++;
  }
  h->ev_used = 0;
  return SOLO_OK;
}

// One physics step = setup -> solve -> finish on `st`.  epilogue != 0 makes the
// finish kernel also run observations / reward / termination.  The solver and the finish
// kernel are programmatic dependents of the kernel in front of them: their blocks are
// scheduled while it drains and wait (cudaGridDependencySynchronize) before they read what
// it wrote.  With per-phase timing on, events sit between the kernels and the launches are
// plain ones.
static int launch_physics(SoloHandle h, const float *actions_dev, int broadcast, float action_scale, int epilogue,
                          float *obs_dev, float *reward_dev, uint8_t *done_dev, float *terminal_obs_dev,
                          cudaStream_t st) {
  const int grid = (h->B.n + h->block - 1) / h->block;
  const size_t sm = tile_smem(h);
  const float margin = h->K.M.margin;
  const bool dbg = h->B.dbg_normal_force || h->B.dbg_motor_torque;
  const int ndof = h->K.M.ndof, n_cand = h->K.M.n_cand;
  int rc;
  cudaEvent_t *ev = nullptr;
  if (h->timing) {
    if (h->ev_used + 4 > h->ev.size()) {
      if ((rc = drain_phase_events(h))) return rc;
    }
    ev = h->ev.data() + h->ev_used;
    h->ev_used += 4;
    cudaEventRecord(ev[0], st);
  }
  const bool pdl = !h->timing;
#define SOLO_MARK(i) do { if (ev) cudaEventRecord(ev[i], st); } while (0)
#define SOLO_PHYSICS(CH)                                                                                          \
  do {                                                                                                            \
    SOLO_SETUP_LAUNCH(CH);                                                                                        \
    if ((rc = check_launch(h, "setup_kernel"))) return rc;                                                        \
    SOLO_MARK(1);                                                                                                 \
    if (dbg) launch_ex(solve_warp_kernel<CH, true, 0, false>, h->B.n, 32, 0, st, pdl, h->K.C, h->B, h->motor_mask, \
                       ndof, n_cand, margin);                                                                     \
    else if (h->solve_mode == 2) launch_ex(solve_warp_kernel<CH, false, 0, true>, persist_grid<CH>(h), 32, 0, st, \
                                           pdl, h->K.C, h->B, h->motor_mask, ndof, n_cand, margin);               \
    else launch_ex(solve_warp_kernel<CH, false, SOLO_MINB0(CH), false>, h->B.n, 32, 0, st, pdl, h->K.C, h->B, h->motor_mask,   \
                   ndof, n_cand, margin);                                                                         \
    if ((rc = check_launch(h, "solve_warp_kernel"))) return rc;                                                   \
    SOLO_MARK(2);                                                                                                 \
    launch_ex(finish_kernel<CH>, grid, h->block, sm, st, pdl, h->K, h->B, epilogue, (int)(pdl && h->warm_finish), \
              obs_dev, reward_dev, done_dev, terminal_obs_dev);                                                   \
  } while (0)
#ifdef SOLO_EXPERIMENTS
#define SOLO_SETUP_LAUNCH(CH)                                                                                     \
  do {                                                                                                            \
    if (!h->legpar) setup_kernel<CH><<<grid, h->block, sm, st>>>(h->K, h->B, actions_dev, broadcast, action_scale); \
    else launch_ex(setup_legpar_kernel<CH>, grid, 4 * h->block, setup_smem(h), st, false, h->K, h->B,             \
                   actions_dev, broadcast, action_scale);                                                                      \
  } while (0)
#else
#define SOLO_SETUP_LAUNCH(CH)                                                                                     \
  launch_ex(setup_legpar_kernel<CH>, grid, 4 * h->block, setup_smem(h), st, false, h->K, h->B, actions_dev,       \
            broadcast, action_scale)
#endif
  if (h->ch == 2) SOLO_PHYSICS(2);
  else SOLO_PHYSICS(3);
  SOLO_MARK(3);
#undef SOLO_SETUP_LAUNCH
#undef SOLO_PHYSICS
#undef SOLO_MARK
  return check_launch(h, "finish_kernel");
}

int solo_simulate(SoloHandle h, const float *actions_dev, int32_t broadcast, int32_t n_steps, void *stream) {
  if (!h || !actions_dev) return fail(h, SOLO_EINVAL, "null argument");
  if (!h->bound) return fail(h, SOLO_EINVAL, "solo_bind has not been called");
  if (n_steps < 0) return fail(h, SOLO_EINVAL, "n_steps must be >= 0");
  DeviceGuard g(h->device);
  for (int k = 0; k < n_steps; k++) {
    int rc = launch_physics(h, actions_dev, broadcast, 1.f, 0, nullptr, nullptr, nullptr, nullptr,
                            (cudaStream_t)stream);
    if (rc) return rc;
  }
  return SOLO_OK;
}

// The launches of one env.step, directly into `st`.
static int step_launches(SoloHandle h, const float *actions_dev, float action_scale, float *obs_dev,
                         float *reward_dev, uint8_t *done_dev, float *terminal_obs_dev, cudaStream_t st) {
  const int sub = h->K.C.substeps;
  for (int k = 0; k < sub; k++) {
    const int last = k == sub - 1;
    int rc = launch_physics(h, actions_dev, 0, action_scale, last, obs_dev, reward_dev, done_dev, terminal_obs_dev, st);
    if (rc) return rc;
  }
  return SOLO_OK;
}

// Capture the step once for this set of output buffers.
static StepGraph *capture_step(SoloHandle h, const float *actions_dev, float action_scale, float *obs_dev,
                               float *reward_dev, uint8_t *done_dev, float *terminal_obs_dev, cudaStream_t st) {
  StepGraph *g = new StepGraph();
  g->graph = nullptr; g->exec = nullptr; g->args = nullptr;
  const int64_t launches0 = h->launches;
  cudaError_t e = cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
  if (e != cudaSuccess) { cudaGetLastError(); delete g; return nullptr; }
  int rc = step_launches(h, actions_dev, action_scale, obs_dev, reward_dev, done_dev, terminal_obs_dev, st);
  e = cudaStreamEndCapture(st, &g->graph);
  h->launches = launches0;     // nothing ran yet: the replay counts them
  if (rc || e != cudaSuccess || !g->graph) {
    cudaGetLastError();
    if (g->graph) cudaGraphDestroy(g->graph);
    delete g;
    return nullptr;
  }
  if (cudaGraphInstantiate(&g->exec, g->graph, 0) != cudaSuccess) {
    cudaGetLastError();
    cudaGraphDestroy(g->graph);
    delete g;
    return nullptr;
  }
  // the setup nodes: their action pointer is the one parameter that changes from call to call
  size_t n_nodes = 0;
  cudaGraphGetNodes(g->graph, nullptr, &n_nodes);
  std::vector<cudaGraphNode_t> nodes(n_nodes);
  cudaGraphGetNodes(g->graph, nodes.data(), &n_nodes);
  const void *setup_fn = h->ch == 2 ? (const void *)setup_legpar_kernel<2> : (const void *)setup_legpar_kernel<3>;
  g->args = new SetupArgs();
  g->args->K = h->K; g->args->B = h->B; g->args->actions = actions_dev; g->args->broadcast = 0;
  g->args->scale = action_scale;
  for (cudaGraphNode_t nd : nodes) {
    cudaGraphNodeType t;
    if (cudaGraphNodeGetType(nd, &t) != cudaSuccess || t != cudaGraphNodeTypeKernel) continue;
    cudaKernelNodeParams kp;
    if (cudaGraphKernelNodeGetParams(nd, &kp) != cudaSuccess) continue;
    if (kp.func != setup_fn) continue;
    kp.kernelParams = g->args->ptrs;
    kp.extra = nullptr;
    g->setup_nodes.push_back(nd);
    g->setup_params.push_back(kp);
  }
  if ((int)g->setup_nodes.size() != h->K.C.substeps) {   // cannot patch what we cannot find: no graph
    cudaGetLastError();
    cudaGraphExecDestroy(g->exec); cudaGraphDestroy(g->graph);
    delete g->args; delete g;
    return nullptr;
  }
  g->actions = actions_dev; g->scale = action_scale;
  g->obs = obs_dev; g->reward = reward_dev; g->done = done_dev; g->terminal = terminal_obs_dev;
  g->epoch = h->epoch; g->used = 0;
  return g;
}

static int step_impl(SoloHandle h, const float *actions_dev, float action_scale, float *obs_dev, float *reward_dev,
                     uint8_t *done_dev, float *terminal_obs_dev, cudaStream_t st, cudaStream_t *ran = nullptr) {
  if (ran) *ran = st;
  if (!h->use_graph || h->timing) return step_launches(h, actions_dev, action_scale, obs_dev, reward_dev, done_dev, terminal_obs_dev, st);
  // retire graphs captured under older bindings / programs
  if (!h->graphs.empty() && h->graphs[0]->epoch != h->epoch) drop_graphs(h);
  // the legacy default stream cannot be captured: use a blocking stream of our own, which the
  // legacy stream orders itself against in both directions
  cudaStream_t run = st;
  if (st == nullptr || st == cudaStreamLegacy) {
    if (!h->gstream && cudaStreamCreate(&h->gstream) != cudaSuccess) {
      cudaGetLastError();
      h->use_graph = false;
      return step_launches(h, actions_dev, action_scale, obs_dev, reward_dev, done_dev, terminal_obs_dev, st);
    }
    run = h->gstream;
  } else if (st == cudaStreamPerThread) {
    return step_launches(h, actions_dev, action_scale, obs_dev, reward_dev, done_dev, terminal_obs_dev, st);
  }
  static uint64_t tick = 0;
  // One graph per output set (env.step hands out a ring of two; the host entry has one pinned
  // block).  The action buffer changes from call to call -- fresh tensors, a pool of pinned host
  // buffers -- and is patched into the setup nodes (2-3 us; capturing and instantiating a graph
  // per buffer costs 1.4 ms each, measured).
  constexpr size_t kMaxGraphs = 8;
  StepGraph *g = nullptr;
  for (StepGraph *c : h->graphs)
    if (c->obs == obs_dev && c->reward == reward_dev && c->done == done_dev && c->terminal == terminal_obs_dev &&
        c->scale == action_scale) { g = c; break; }
  if (!g) {
    if (h->graphs.size() >= kMaxGraphs) {   // callers that allocate fresh outputs every step: evict the oldest
      size_t lru = 0;
      for (size_t i = 1; i < h->graphs.size(); i++) if (h->graphs[i]->used < h->graphs[lru]->used) lru = i;
      StepGraph *old = h->graphs[lru];
      cudaGraphExecDestroy(old->exec); cudaGraphDestroy(old->graph); delete old->args; delete old;
      h->graphs.erase(h->graphs.begin() + lru);
    }
    g = capture_step(h, actions_dev, action_scale, obs_dev, reward_dev, done_dev, terminal_obs_dev, run);
    if (!g) {   // capture refused (another capture in flight, ...): plain launches, and stop trying
      h->use_graph = false;
      return step_launches(h, actions_dev, action_scale, obs_dev, reward_dev, done_dev, terminal_obs_dev, st);
    }
    h->graphs.push_back(g);
  }
  if (g->actions != actions_dev) {
    g->args->actions = actions_dev;
    for (size_t i = 0; i < g->setup_nodes.size(); i++) {
      cudaError_t e = cudaGraphExecKernelNodeSetParams(g->exec, g->setup_nodes[i], &g->setup_params[i]);
      if (e != cudaSuccess) return cuda_fail(h, e, "cudaGraphExecKernelNodeSetParams");
    }
    g->actions = actions_dev;
  }
  g->used = ++tick;
  if (ran) *ran = run;
  cudaError_t e = cudaGraphLaunch(g->exec, run);
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaGraphLaunch");
  h->launches += 3 * (int64_t)h->K.C.substeps;
  return SOLO_OK;
}

int solo_step(SoloHandle h, const float *actions_dev, float action_scale, float *obs_dev, float *reward_dev,
              uint8_t *done_dev, float *terminal_obs_dev, void *stream) {
  if (!h || !actions_dev || !obs_dev || !reward_dev || !done_dev) return fail(h, SOLO_EINVAL, "null argument");
  if (!h->bound) return fail(h, SOLO_EINVAL, "solo_bind has not been called");
  if (!h->program_set || h->obs_dim < 1) return fail(h, SOLO_EINVAL, "no observation program set");
  DeviceGuard g(h->device);
  return step_impl(h, actions_dev, action_scale, obs_dev, reward_dev, done_dev, terminal_obs_dev, (cudaStream_t)stream);
}

static int observe_impl(SoloHandle h, float *obs_dev, float *reward_dev, void *stream) {
  if (!h->bound) return fail(h, SOLO_EINVAL, "solo_bind has not been called");
  if (!h->program_set) return fail(h, SOLO_EINVAL, "no program set");
  if (obs_dev && h->obs_dim < 1) return fail(h, SOLO_EINVAL, "no observation program set");
  DeviceGuard g(h->device);
  const int grid = (h->B.n + h->block - 1) / h->block;
  const size_t sm = tile_smem(h);
  cudaStream_t st = (cudaStream_t)stream;
  if (h->ch == 2) observe_kernel<2><<<grid, h->block, sm, st>>>(h->K, h->B, obs_dev, reward_dev);
  else observe_kernel<3><<<grid, h->block, sm, st>>>(h->K, h->B, obs_dev, reward_dev);
  return check_launch(h, "observe_kernel");
}

int solo_observe(SoloHandle h, float *obs_dev, void *stream) {
  if (!h || !obs_dev) return fail(h, SOLO_EINVAL, "null argument");
  return observe_impl(h, obs_dev, nullptr, stream);
}

int solo_reward(SoloHandle h, float *reward_dev, void *stream) {
  if (!h || !reward_dev) return fail(h, SOLO_EINVAL, "null argument");
  if (h->program_set && h->K.P.n_rew < 1) return fail(h, SOLO_EINVAL, "no reward program set");
  return observe_impl(h, nullptr, reward_dev, stream);
}

int solo_reset(SoloHandle h, const uint8_t *mask_dev, void *stream) {
  if (!h) return SOLO_EINVAL;
  if (!h->bound) return fail(h, SOLO_EINVAL, "solo_bind has not been called");
  DeviceGuard g(h->device);
  const int block = 128;
  const int grid = (h->B.n + block - 1) / block;
  reset_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(h->B, h->state_dim, mask_dev);
  return check_launch(h, "reset_kernel");
}

// Device alias of a host pointer the GPU can reach directly (page-locked memory under unified
// addressing: cudaHostAlloc / cudaHostRegister, torch's pin_memory()), or null.
static void *device_alias(const void *host) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}

int solo_step_host(SoloHandle h, const float *actions_host, float action_scale, float *obs_host,
                   float *reward_host, uint8_t *done_host, void *stream) {
  if (!h || !actions_host || !obs_host || !reward_host || !done_host) return fail(h, SOLO_EINVAL, "null argument");
  if (!h->bound) return fail(h, SOLO_EINVAL, "solo_bind has not been called");
  if (!h->program_set || h->obs_dim < 1) return fail(h, SOLO_EINVAL, "no observation program set");
  DeviceGuard g(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n = (size_t)h->B.n;
  // Page-locked caller buffers: the kernels read the actions and write observations / reward /
  // done THROUGH the caller's own memory (the bytes cross the bus inside the step, overlapped
  // with the kernels, instead of as separate copies in front of and behind it).  The aliases of
  // the three result buffers are looked up when the triple changes, the action buffer's on
  // every call.
  const void *keys[3] = {obs_host, reward_host, done_host};
  if (keys[0] != h->map_key[0] || keys[1] != h->map_key[1] || keys[2] != h->map_key[2]) {
    for (int i = 0; i < 3; i++) { h->map_key[i] = keys[i]; h->map_dev[i] = device_alias(keys[i]); }
  }
  void *act_dev = device_alias(actions_host);
  if (h->zero_copy && act_dev && h->map_dev[0] && h->map_dev[1] && h->map_dev[2]) {
    cudaStream_t ran = st;
    int rc = step_impl(h, (const float *)act_dev, action_scale, (float *)h->map_dev[0], (float *)h->map_dev[1],
                       (uint8_t *)h->map_dev[2], nullptr, st, &ran);
    if (rc) return rc;
    // the stream the step actually ran on (our own blocking stream when the caller's is the legacy one)
    cudaError_t e = cudaStreamSynchronize(ran);
    if (e != cudaSuccess) return cuda_fail(h, e, "step (asynchronous fault in setup / solve / finish)");
    return SOLO_OK;
  }
  // pageable buffers: through the bound staging buffers
  if (!h->st_actions) return fail(h, SOLO_EINVAL, "solo_bind_staging has not been called (needed for pageable host buffers)");
  cudaError_t e = cudaMemcpyAsync(h->st_actions, actions_host, n * h->K.M.n_api * sizeof(float),
                                  cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return cuda_fail(h, e, "H2D actions");
  int rc = step_launches(h, h->st_actions, action_scale, h->st_obs, h->st_reward, h->st_done, nullptr, st);
  if (rc) return rc;
  const size_t ob = n * h->obs_dim * sizeof(float), rb = n * sizeof(float);
  const bool packed = (char *)h->st_reward == (char *)h->st_obs + ob && (char *)h->st_done == (char *)h->st_reward + rb &&
                      (char *)reward_host == (char *)obs_host + ob && (char *)done_host == (char *)reward_host + rb;
  if (packed) {
    // staging and host buffers are both laid out obs | reward | done: one copy
    e = cudaMemcpyAsync(obs_host, h->st_obs, ob + rb + n, cudaMemcpyDeviceToHost, st);
  } else {
    e = cudaMemcpyAsync(obs_host, h->st_obs, ob, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(reward_host, h->st_reward, rb, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(done_host, h->st_done, n, cudaMemcpyDeviceToHost, st);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return cuda_fail(h, e, "D2H results");
  return SOLO_OK;
}

int solo_set_option(SoloHandle h, int32_t option, int32_t value) {
  if (!h) return SOLO_EINVAL;
  DeviceGuard g(h->device);
  switch (option) {
    case SOLO_OPT_GRAPH:
      h->use_graph = value != 0;
      if (!h->use_graph) drop_graphs(h);
      return SOLO_OK;
    case SOLO_OPT_ZERO_COPY:
      h->zero_copy = value != 0;
      return SOLO_OK;
    case SOLO_OPT_FINISH_WARMUP:
      h->warm_finish = value < 0 ? h->B.n < kHiOccEnvs : value != 0;
      drop_graphs(h);
      return SOLO_OK;
    case SOLO_OPT_SOLVER_OCCUPANCY:
      if (value > 2) return fail(h, SOLO_EINVAL, "SOLO_OPT_SOLVER_OCCUPANCY: 0, 1, 2 or -1");
      h->solve_mode = value < 0 ? auto_solve_mode(h->B.n) : value;
      drop_graphs(h);
      return SOLO_OK;
    default:
      return fail(h, SOLO_EINVAL, "unknown option");
  }
}

int64_t solo_launch_count(SoloHandle h) { return h ? h->launches : 0; }

int solo_phase_timing(SoloHandle h, int32_t enable) {
  if (!h) return SOLO_EINVAL;
  DeviceGuard g(h->device);
  if (enable && h->ev.empty()) {
    h->ev.resize(4 * 1024);
    for (auto &e : h->ev)
      if (cudaEventCreate(&e) != cudaSuccess) return fail(h, SOLO_ECUDA, "cudaEventCreate failed");
  }
  h->timing = enable != 0;
  h->ev_used = 0;
  h->phase_ms[0] = h->phase_ms[1] = h->phase_ms[2] = 0;
  h->phase_steps = 0;
  return SOLO_OK;
}

int64_t solo_phase_times(SoloHandle h, double *ms3) {
  if (!h || !ms3) return SOLO_EINVAL;
  DeviceGuard g(h->device);
  if (drain_phase_events(h)) return SOLO_ECUDA;
  for (int p = 0; p < 3; p++) ms3[p] = h->phase_ms[p];
  return h->phase_steps;
}

int solo_block_size(SoloHandle h) { return h ? h->block : SOLO_EINVAL; }

int solo_solver_stats(SoloHandle h, const float **stats_dev) {
  if (!h || !stats_dev) return SOLO_EINVAL;
  if (!h->bound) return fail(h, SOLO_EINVAL, "solo_bind has not been called");
  *stats_dev = h->B.yh + (size_t)(6 + kLegs * h->ch) * (size_t)h->B.n;
  return SOLO_OK;
}

// ---- render() ------------------------------------------------------------------------
int solo_bind_mesh(SoloHandle h, const int32_t *tri_body_dev, const float *tri_verts_dev,
                   const float *tri_rgba_dev, int32_t n_tri) {
  if (!h) return SOLO_EINVAL;
  if (n_tri < 0 || (n_tri > 0 && (!tri_body_dev || !tri_verts_dev || !tri_rgba_dev)))
    return fail(h, SOLO_EINVAL, "solo_bind_mesh: null triangle table");
  h->tri_body = tri_body_dev;
  h->tri_verts = tri_verts_dev;
  h->tri_rgba = tri_rgba_dev;
  h->n_tri = n_tri;
  return SOLO_OK;
}

static size_t render_recs_bytes(const SoloHandle h, int64_t env_count) {
  return align256(sizeof(TriRec) * (size_t)(h->n_tri > 0 ? h->n_tri : 1) * (size_t)env_count);
}

int64_t solo_render_scratch_bytes(SoloHandle h, int32_t env_count) {
  if (!h || env_count < 1) return SOLO_EINVAL;
  return (int64_t)(render_recs_bytes(h, env_count) + align256(sizeof(int32_t) * 4 * (size_t)env_count));
}

int solo_render(SoloHandle h, const SoloCamera *cam, int32_t env_begin, int32_t env_count, uint8_t *rgba_dev,
                float *depth_dev, int32_t *seg_dev, void *scratch_dev, void *stream) {
  if (!h || !cam || !rgba_dev || !scratch_dev) return fail(h, SOLO_EINVAL, "null argument");
  if (!h->bound) return fail(h, SOLO_EINVAL, "solo_bind has not been called");
  if (env_begin < 0 || env_count < 1 || (int64_t)env_begin + env_count > h->B.n)
    return fail(h, SOLO_EINVAL, "solo_render: environment range out of bounds");
  if (env_count > 65535) return fail(h, SOLO_EINVAL, "solo_render: at most 65535 environments per call");
  if (cam->width < 1 || cam->height < 1 || cam->width > 16384 || cam->height > 16384)
    return fail(h, SOLO_EINVAL, "solo_render: bad image size");
  if (((uintptr_t)scratch_dev & 255) != 0 || ((uintptr_t)rgba_dev & 3) != 0)
    return fail(h, SOLO_EINVAL, "solo_render: scratch must be 256-byte aligned, the image 4-byte aligned");
  RenderParams R;
  std::string msg = make_render_params(*cam, &R);
  if (!msg.empty()) return fail(h, SOLO_EINVAL, msg);
  R.n_tri = h->n_tri;
  R.env_begin = env_begin;
  R.env_count = env_count;
  // one frame alone needs every block it can get (32-pixel tiles, one pixel per thread); a batch of
  // frames fills the GPU anyway (128-pixel tiles, four pixels per thread, four tiles per block)
  const int pix = env_count >= 32 ? 4 : 1;
  const int tiles_x = (cam->width + kTileW * pix - 1) / (kTileW * pix);
  const int tiles_y = (cam->height + kTileH - 1) / kTileH;
  DeviceGuard g(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  TriRec *recs = (TriRec *)scratch_dev;
  int32_t *bbox = (int32_t *)((char *)scratch_dev + render_recs_bytes(h, env_count));
  cudaError_t e = cudaMemsetAsync(bbox, 0x7f, sizeof(int32_t) * 4 * (size_t)env_count, st);
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaMemsetAsync");
  if (h->n_tri > 0) {
    const dim3 grid((h->n_tri + kProjectThreads - 1) / kProjectThreads, env_count);
    render_project_kernel<<<grid, kProjectThreads, 0, st>>>(h->K.M, R, h->B.state, h->B.n, h->tri_body,
                                                             h->tri_verts, h->tri_rgba, recs, bbox);
    int rc = check_launch(h, "render_project_kernel");
    if (rc) return rc;
  }
  R.tiles_per_block = env_count >= 32 ? 4 : 1;
  const dim3 grid(tiles_x, (tiles_y + R.tiles_per_block - 1) / R.tiles_per_block, env_count);
  const bool aux = depth_dev || seg_dev;
#define SOLO_RASTER(P, A) render_raster_kernel<P, A><<<grid, kRasterThreads, 0, st>>>(R, recs, bbox, (uint32_t *)rgba_dev, depth_dev, seg_dev)
  if (pix == 4) { if (aux) SOLO_RASTER(4, true); else SOLO_RASTER(4, false); }
  else { if (aux) SOLO_RASTER(1, true); else SOLO_RASTER(1, false); }
#undef SOLO_RASTER
  return check_launch(h, "render_raster_kernel");
}

}  // extern "C"

// ---- solo_render_static: render() for callers that cannot allocate device memory ----
// The projected-triangle records and bounding boxes of one call live in this file-scope array
// instead of caller scratch.  In host code `g_render_static` is only a stand-in for the symbol:
// its device address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_render_static[SOLO_RENDER_STATIC_BYTES];

namespace {
constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mutex;
void *g_static_ptr[kMaxStaticDevices] = {};
}  // namespace

extern "C" int solo_render_static(SoloHandle h, const SoloCamera *cam, int32_t env_begin, int32_t env_count,
                                  uint8_t *rgba_dev, float *depth_dev, int32_t *seg_dev, void *stream) {
  if (!h || !cam || !rgba_dev) return fail(h, SOLO_EINVAL, "null argument");
  if (env_count < 1) return fail(h, SOLO_EINVAL, "solo_render_static: environment range out of bounds");
  const int64_t need = solo_render_scratch_bytes(h, env_count);
  if (need > (int64_t)SOLO_RENDER_STATIC_BYTES) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "solo_render_static: %d environments x %d triangles need %lld bytes of scratch, the static array holds "
             "%lld (SOLO_RENDER_STATIC_BYTES); render fewer environments per call or use solo_render with caller scratch",
             (int)env_count, (int)h->n_tri, (long long)need, (long long)SOLO_RENDER_STATIC_BYTES);
    return fail(h, SOLO_EINVAL, msg);
  }
  if (h->device < 0 || h->device >= kMaxStaticDevices) return fail(h, SOLO_EINVAL, "solo_render_static: device index out of range");
  void *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    scratch = g_static_ptr[h->device];
    if (!scratch) {
      DeviceGuard g(h->device);
      if (!g.ok) return fail(h, SOLO_ECUDA, "cudaSetDevice failed");
      const cudaError_t e = cudaGetSymbolAddress(&scratch, g_render_static);
      if (e != cudaSuccess || !scratch) return cuda_fail(h, e, "cudaGetSymbolAddress(g_render_static)");
      g_static_ptr[h->device] = scratch;
    }
  }
  // the array keeps the previous call's records: solo_render resets the bounding boxes (the only
  // part read before it is written) with a memset on `stream`; every record read is rewritten first
  return solo_render(h, cam, env_begin, env_count, rgba_dev, depth_dev, seg_dev, scratch, stream);
}
