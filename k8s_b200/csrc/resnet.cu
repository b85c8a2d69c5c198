// ResNet-50 (v1.5: stride on the 3x3 of each bottleneck) forward + backward for one local
// replica -- BASELINE config 4.  bf16 activations and operand weights,
// fp32 master weights / statistics / gradients of parameters.
//
// Every convolution is a tcgen05 GEMM (gemm_tc.cuh): 1x1 stride-1 convolutions read the NHWC
// activation as the [pixels, C] operand in place; the others go through an explicit bf16
// im2col (`col`, one buffer per layer, kept from forward for the weight gradient: 7 GB at batch 256).  For a conv with input x, filter W
// (HWIO) and output gradient dz:
//     forward  z[M, Kout]     = col[M, RSC] (K-major)  x W[RSC, Kout] (MN-major)  -> bf16
//     wgrad    dW[RSC, Kout]  = col[M, RSC] (MN-major) x dz[M, Kout]  (MN-major)  -> fp32, split-K
//     dgrad    dcol[M, RSC]   = dz[M, Kout] (K-major)  x W[RSC, Kout] (K-major)   -> bf16, then col2im
// Batch norm uses batch statistics (two-stage fixed-order reductions), eps 1e-5, and keeps
// moving statistics with decay 0.997.
#include <algorithm>
#include <array>
#include <cmath>
#include <string>
#include <vector>

#include "common.cuh"
#include "resnet_kernels.cuh"
#include "conv_host.cuh"
#include "tc_host.cuh"

namespace tfg {

using rn::bf16;

struct RnTensor {
  std::string name;
  long long n = 0, h = 0, w = 0, c = 0;
  long long elems() const { return n * h * w * c; }
};

struct RnUnit {                      // conv + batch norm (+ residual add) (+ relu)
  std::string name;
  rn::ConvGeom g;
  int kout = 0;
  int in = -1, res = -1, z = -1, y = -1;   // tensor ids
  bool relu = true, direct = false;        // direct: 1x1 stride 1, the activation itself is the GEMM operand
  bool halo = false;                       // 3x3 stride 1 with the input tile resident in shared memory (conv3x3_halo_kernel)
  bool implicit = false;                   // implicit GEMM through 4-D TMA boxes (C % 64 == 0): no im2col either
  int v_w = -1, v_gamma = -1, v_beta = -1;
  float *mov_mean = nullptr, *mov_var = nullptr;   // per job
};

struct RnSlot {
  float* x_stage = nullptr;          // [B,H,W,3] fp32 staging for host inputs
  int32_t* lab_stage = nullptr;
  std::vector<bf16*> data, grad;     // per tensor
  std::vector<float*> mean, invstd;  // per unit
  std::vector<bf16*> col;            // per unit: im2col of its input, kept from forward to backward (NULL if direct)
  bf16* dcol = nullptr;              // shared staging of data gradients
  uint8_t* pool_arg = nullptr;
  float *bn_partial = nullptr, *splitk = nullptr;
  float* tile_stats = nullptr;             // per GEMM tile column sums of z (fused batch-norm statistics)
  std::vector<uint8_t*> relu_bits;         // per relu unit: one bit per element of y, set where the relu passed
  unsigned int* bn_done = nullptr;         // per 64-channel group: row chunks finished (self-resetting)
  float *logits = nullptr, *loss_vec = nullptr, *dbn = nullptr;   // dbn: [2][Cmax] dbeta, dgamma staging unused
  bf16 *pooled = nullptr, *dpooled = nullptr, *dlogits = nullptr;
  std::vector<TcGemm> fwd, wgrad, dgrad;   // per unit (explicit-operand GEMMs)
  std::vector<HaloConv> hfwd, hdgrad;      // per unit (halo kernel)
  std::vector<std::array<ConvGemm, 4>> s2dgrad;   // per stride-2 implicit unit: one data-gradient GEMM per output parity class (ntaps == 0: none)
  std::vector<ConvGemm> ifwd, idgrad;      // per unit (implicit GEMM)
  std::vector<ConvWgrad> iwgrad;
  TcGemm fc_fwd, fc_wgrad, fc_dgrad;
  bool maps_ready = false;
};

struct ResnetModel {
  int B = 0, HW = 0, classes = 0;
  std::vector<RnTensor> tensors;
  std::vector<RnUnit> units;
  int t_image = -1, t_pool_in = -1, t_pool_out = -1, t_last = -1;
  int v_fcw = -1, v_fcb = -1;
  size_t col_elems = 0, splitk_floats = 0, tile_stats_floats = 0;
  int cmax = 0;
  std::vector<RnSlot> slots;
};

static inline int grid_for(long long n, int per = 256) {
  long long b = (n + per - 1) / per;
  if (b > kNumSMs * 16) b = kNumSMs * 16;
  return (int)std::max<long long>(b, 1);
}

// row chunks of the batch-norm reductions: about four resident blocks per SM over all channel
// groups, at least 32 rows each
constexpr int kBnTargetBlocks = kNumSMs * 4;
constexpr size_t kBnPartialFloats = (size_t)(kBnTargetBlocks + 32) * 64;   // chunks * C never exceeds this

// bf16 GEMM outputs leave through TMA stores / reduce-adds wherever the shape allows it (measured against
// register-direct stores plus a separate accumulate kernel: profiles/README.md, round 1)
static constexpr bool tma_out_enabled() { return true; }

// element-wise kernels with a grid-stride loop and three resident blocks per SM: one full wave
static inline int grid_resident(long long n, int per = 256) {
  return (int)std::max<long long>(1, std::min<long long>((n + per - 1) / per, kNumSMs * 3));
}

// stride-2 data gradients run as one implicit GEMM per parity class (measured against dz x W^T plus a
// col2im gather: -1.6 ms per step, profiles/README.md, round 1)
static constexpr bool s2_dgrad_enabled() { return true; }

#define RN_OK(j, what)                                           \
  do {                                                           \
    prof_mark((j), what);                                        \
    cudaError_t e__ = cudaGetLastError();                        \
    if (e__ != cudaSuccess) return fail_cuda((j), e__, what);    \
  } while (0)

// ---- model construction ----------------------------------------------------------------------
static int add_tensor(ResnetModel* m, const std::string& name, long long n, long long h, long long w, long long c) {
  RnTensor t;
  t.name = name; t.n = n; t.h = h; t.w = w; t.c = c;
  m->tensors.push_back(t);
  return (int)m->tensors.size() - 1;
}

static int reg(tfg_job* j, const std::string& name, std::vector<int64_t> shape, bool trainable, int* var) {
  return tfg_var_register(j, name.c_str(), "VariableV2", TFG_DT_F32, shape.data(), (int)shape.size(), trainable ? 1 : 0,
                          var, nullptr);
}

// split-K factor of a weight-gradient GEMM with `tiles` output tiles: as many splits as keep
// tiles * splits within ONE wave of CTAs (rounding up instead leaves a second wave of a few CTAs
// that doubles the kernel's time)
static long long wgrad_splits(long long tiles) {
  return std::max<long long>(1, std::min<long long>(64, kNumSMs / tiles));
}

// number of [2][kout] column-sum records the forward GEMM of a unit leaves for the batch-norm fold:
// one per CTA row of the persistent grid and row slice (tc::store_col_stats)
static int stat_tiles(const RnUnit& u) {
  const int bn = u.kout >= 128 ? 128 : 64;
  const int parts = bn == 128 ? tc::col_stats_parts<128>() : tc::col_stats_parts<64>();
  const int tiles_n = (u.kout + bn - 1) / bn;
  long long m_tiles;
  if (u.halo) {
    tc::HaloShape hs;
    fill_halo_shape(&hs, u.g.N, u.g.OH, u.g.OW, u.g.C, 0, tiles_n);
    return parts * halo_grid_rows(hs, tiles_n);
  }
  if (u.implicit) {
    int tn, th;
    pick_conv_tile(u.g.N, u.g.OH, u.g.OW, false, &tn, &th);
    m_tiles = (long long)((u.g.N + tn - 1) / tn) * ((u.g.OH + th - 1) / th);
  } else {
    m_tiles = ((long long)u.g.N * u.g.OH * u.g.OW + 127) / 128;
  }
  return parts * (tc::persistent_grid(m_tiles * tiles_n, tiles_n, true, kNumSMs) / tiles_n);
}

// conv(R, stride, kout) + BN on tensor `in`; returns the unit index
static int add_unit(tfg_job* j, ResnetModel* m, const std::string& name, int in, int R, int stride, int kout, bool relu,
                    int res, int* rc) {
  const RnTensor x = m->tensors[in];
  if (kout % 64 != 0 || 2048 % kout != 0) {   // the batch-norm kernels keep 8 fixed channels per thread
    *rc = fail(j, TFG_EINVAL, "internal: " + name + " has " + std::to_string(kout) + " output channels; need a divisor of 2048 >= 64");
    return -1;
  }
  RnUnit u;
  u.name = name;
  u.g.N = (int)x.n; u.g.H = (int)x.h; u.g.W = (int)x.w; u.g.C = (int)x.c;
  u.g.R = u.g.S = R; u.g.stride = stride; u.g.pad = R / 2;
  u.g.OH = (u.g.H + 2 * u.g.pad - R) / stride + 1;
  u.g.OW = (u.g.W + 2 * u.g.pad - R) / stride + 1;
  u.g.Kp = (R * R * u.g.C + 7) / 8 * 8;
  u.kout = kout; u.in = in; u.res = res; u.relu = relu;
  u.direct = (R == 1 && stride == 1);
  u.implicit = !u.direct && (x.c % 64 == 0);
  u.halo = u.implicit && tma_out_enabled() && halo_fits(u.g.OH, u.g.OW, u.g.C, R, stride);
  u.z = add_tensor(m, name + "/z", x.n, u.g.OH, u.g.OW, kout);
  u.y = add_tensor(m, name + "/y", x.n, u.g.OH, u.g.OW, kout);
  int dummy;
  if ((*rc = reg(j, name + "/weights", {R, R, x.c, kout}, true, &u.v_w))) return -1;
  if ((*rc = reg(j, name + "/bn/gamma", {kout}, true, &u.v_gamma))) return -1;
  if ((*rc = reg(j, name + "/bn/beta", {kout}, true, &u.v_beta))) return -1;
  if ((*rc = reg(j, name + "/bn/moving_mean", {kout}, false, &dummy))) return -1;
  if ((*rc = reg(j, name + "/bn/moving_variance", {kout}, false, &dummy))) return -1;
  const long long M = x.n * u.g.OH * u.g.OW;
  if (!u.direct) m->col_elems = std::max<size_t>(m->col_elems, (size_t)M * u.g.Kp);
  else m->col_elems = std::max<size_t>(m->col_elems, (size_t)M * x.c);   // dgrad staging when accumulating
  m->cmax = std::max(m->cmax, kout);
  m->units.push_back(u);
  return (int)m->units.size() - 1;
}

int resnet_build(tfg_job* j, int batch, int image_hw, int classes) {
  ResnetModel* m = new ResnetModel();
  m->B = batch; m->HW = image_hw; m->classes = classes;
  int rc = TFG_OK;
  m->t_image = add_tensor(m, "image", batch, image_hw, image_hw, 3);
  int u0 = add_unit(j, m, "conv1", m->t_image, 7, 2, 64, true, -1, &rc);
  if (u0 < 0) { delete m; return rc; }
  m->t_pool_in = m->units[u0].y;
  const RnTensor y0 = m->tensors[m->t_pool_in];
  m->t_pool_out = add_tensor(m, "pool1", batch, (y0.h + 2 - 3) / 2 + 1, (y0.w + 2 - 3) / 2 + 1, 64);
  int cur = m->t_pool_out;
  const int blocks[4] = {3, 4, 6, 3}, mids[4] = {64, 128, 256, 512};
  for (int st = 0; st < 4; ++st)
    for (int b = 0; b < blocks[st]; ++b) {
      const std::string base = "stage" + std::to_string(st + 1) + "/block" + std::to_string(b + 1);
      const int stride = (b == 0 && st > 0) ? 2 : 1;
      const int mid = mids[st], out = 4 * mid;
      int shortcut = cur;
      if (b == 0) {   // projection shortcut
        int us = add_unit(j, m, base + "/shortcut", cur, 1, stride, out, false, -1, &rc);
        if (us < 0) { delete m; return rc; }
        shortcut = m->units[us].y;
      }
      int ua = add_unit(j, m, base + "/a", cur, 1, 1, mid, true, -1, &rc);
      if (ua < 0) { delete m; return rc; }
      int ub = add_unit(j, m, base + "/b", m->units[ua].y, 3, stride, mid, true, -1, &rc);
      if (ub < 0) { delete m; return rc; }
      int uc = add_unit(j, m, base + "/c", m->units[ub].y, 1, 1, out, true, shortcut, &rc);
      if (uc < 0) { delete m; return rc; }
      cur = m->units[uc].y;
    }
  m->t_last = cur;
  if ((rc = reg(j, "fc/weights", {2048, classes}, true, &m->v_fcw)) || (rc = reg(j, "fc/biases", {classes}, true, &m->v_fcb))) {
    delete m;
    return rc;
  }
  int gs;
  if ((rc = tfg_var_register(j, "global_step", "VariableV2", TFG_DT_I64, nullptr, 0, 0, &gs, nullptr))) { delete m; return rc; }
  // split-K partials: the largest (splits * RSC * Kout) any weight-gradient GEMM may need
  for (auto& u : m->units) {
    const long long rsc = (long long)u.g.R * u.g.S * u.g.C;
    const int bn = u.kout >= 128 ? 128 : 64;
    const long long tiles = ((rsc + 127) / 128) * ((u.kout + bn - 1) / bn);
    const long long splits = wgrad_splits(tiles);
    m->splitk_floats = std::max<size_t>(m->splitk_floats, (size_t)(splits * rsc * u.kout));
    m->tile_stats_floats = std::max<size_t>(m->tile_stats_floats, (size_t)std::max(stat_tiles(u), kNumSMs * 3) * 2 * u.kout);
  }
  m->splitk_floats = std::max<size_t>(m->splitk_floats, (size_t)8 * 2048 * classes);

  // device buffers
  cudaError_t e = cudaSuccess;
  auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes ? bytes : 16); };
  for (auto& u : m->units) {
    alloc((void**)&u.mov_mean, sizeof(float) * u.kout);
    alloc((void**)&u.mov_var, sizeof(float) * u.kout);
  }
  m->slots.resize(j->local_replicas.size());
  for (auto& s : m->slots) {
    alloc((void**)&s.x_stage, sizeof(float) * (size_t)batch * image_hw * image_hw * 3);
    alloc((void**)&s.lab_stage, sizeof(int32_t) * batch);
    s.data.assign(m->tensors.size(), nullptr);
    s.grad.assign(m->tensors.size(), nullptr);
    for (size_t t = 0; t < m->tensors.size(); ++t) {
      alloc((void**)&s.data[t], sizeof(bf16) * (size_t)m->tensors[t].elems());
      if ((int)t != m->t_image) alloc((void**)&s.grad[t], sizeof(bf16) * (size_t)m->tensors[t].elems());
    }
    s.mean.assign(m->units.size(), nullptr);
    s.invstd.assign(m->units.size(), nullptr);
    for (size_t u = 0; u < m->units.size(); ++u) {
      alloc((void**)&s.mean[u], sizeof(float) * m->units[u].kout);
      alloc((void**)&s.invstd[u], sizeof(float) * m->units[u].kout);
    }
    s.relu_bits.assign(m->units.size(), nullptr);
    for (size_t u = 0; u < m->units.size(); ++u) {
      const RnUnit& un = m->units[u];
      if (un.relu) alloc((void**)&s.relu_bits[u], (size_t)un.g.N * un.g.OH * un.g.OW * un.kout / 8);
    }
    s.col.assign(m->units.size(), nullptr);
    for (size_t u = 0; u < m->units.size(); ++u) {
      const RnUnit& un = m->units[u];
      if (!un.direct && !un.implicit)
        alloc((void**)&s.col[u], sizeof(bf16) * (size_t)un.g.N * un.g.OH * un.g.OW * un.g.Kp);
    }
    alloc((void**)&s.dcol, sizeof(bf16) * m->col_elems);
    alloc((void**)&s.pool_arg, (size_t)m->tensors[m->t_pool_out].elems());
    alloc((void**)&s.bn_partial, sizeof(float) * 2 * kBnPartialFloats);
    alloc((void**)&s.tile_stats, sizeof(float) * m->tile_stats_floats);
    alloc((void**)&s.bn_done, sizeof(unsigned int) * 64);
    if (e == cudaSuccess) e = cudaMemset(s.bn_done, 0, sizeof(unsigned int) * 64);
    alloc((void**)&s.splitk, sizeof(float) * m->splitk_floats);
    alloc((void**)&s.logits, sizeof(float) * (size_t)batch * classes);
    alloc((void**)&s.loss_vec, sizeof(float) * batch);
    alloc((void**)&s.pooled, sizeof(bf16) * (size_t)batch * 2048);
    alloc((void**)&s.dpooled, sizeof(bf16) * (size_t)batch * 2048);
    alloc((void**)&s.dlogits, sizeof(bf16) * (size_t)batch * classes);
  }
  j->resnet = m;
  if (e != cudaSuccess) { resnet_destroy(j); return fail_cuda(j, e, "cudaMalloc(ResNet-50 buffers)"); }
  for (auto& u : m->units) {   // moving_mean = 0, moving_variance = 1
    cudaMemsetAsync(u.mov_mean, 0, sizeof(float) * u.kout, j->stream);
    std::vector<float> ones(u.kout, 1.f);
    cudaMemcpyAsync(u.mov_var, ones.data(), sizeof(float) * u.kout, cudaMemcpyHostToDevice, j->stream);
    cudaStreamSynchronize(j->stream);
  }
  return TFG_OK;
}

void resnet_destroy(tfg_job* j) {
  ResnetModel* m = j->resnet;
  if (!m) return;
  for (auto& u : m->units) { cudaFree(u.mov_mean); cudaFree(u.mov_var); }
  for (auto& s : m->slots) {
    cudaFree(s.x_stage); cudaFree(s.lab_stage);
    for (auto p : s.data) cudaFree(p);
    for (auto p : s.grad) cudaFree(p);
    for (auto p : s.mean) cudaFree(p);
    for (auto p : s.invstd) cudaFree(p);
    for (auto p : s.col) cudaFree(p);
    for (auto p : s.relu_bits) cudaFree(p);
    cudaFree(s.dcol); cudaFree(s.pool_arg); cudaFree(s.bn_partial); cudaFree(s.tile_stats); cudaFree(s.bn_done); cudaFree(s.splitk);
    cudaFree(s.logits); cudaFree(s.loss_vec); cudaFree(s.pooled); cudaFree(s.dpooled); cudaFree(s.dlogits);
  }
  delete m;
  j->resnet = nullptr;
}

int resnet_batch(const tfg_job* j) { return j->resnet ? j->resnet->B : 0; }
size_t resnet_image_floats(const tfg_job* j) { return j->resnet ? (size_t)j->resnet->HW * j->resnet->HW * 3 : 0; }
float* resnet_stage_images(tfg_job* j, int slot) { return j->resnet->slots[slot].x_stage; }
int32_t* resnet_stage_labels(tfg_job* j, int slot) { return j->resnet->slots[slot].lab_stage; }

// ---- GEMM descriptors (encoded once per slot, after tfg_vars_finalize) --------------------------
static int prepare_maps(tfg_job* j, int slot) {
  ResnetModel* m = j->resnet;
  RnSlot& s = m->slots[slot];
  if (s.maps_ready) return TFG_OK;
  if (!j->params_lp) return fail(j, TFG_ESTATE, "ResNet-50 needs the bf16 parameter shadow");
  const int kind = TFG_MATH_BF16;
  auto wlp = [&](int v) { return static_cast<const void*>(static_cast<const char*>(j->params_lp) + 2 * j->vars[v].offset); };
  auto grd = [&](int v) { return j->grads[slot] + j->vars[v].offset; };
  const size_t nu = m->units.size();
  s.fwd.assign(nu, TcGemm());
  s.wgrad.assign(nu, TcGemm());
  s.dgrad.assign(nu, TcGemm());
  s.hfwd.assign(nu, HaloConv());
  s.hdgrad.assign(nu, HaloConv());
  s.s2dgrad.assign(nu, std::array<ConvGemm, 4>());
  s.ifwd.assign(nu, ConvGemm());
  s.idgrad.assign(nu, ConvGemm());
  s.iwgrad.assign(nu, ConvWgrad());
  int rc;
  for (size_t i = 0; i < nu; ++i) {
    const RnUnit& u = m->units[i];
    const long long M = (long long)u.g.N * u.g.OH * u.g.OW;
    const int rsc = u.g.R * u.g.S * u.g.C;
    if (u.halo) {
      const int bn = u.kout >= 128 ? 128 : 64, bnd = u.g.C >= 128 ? 128 : 64;
      HaloConv& f = s.hfwd[i];
      fill_halo_shape(&f.hs, u.g.N, u.g.OH, u.g.OW, u.g.C, 0, (u.kout + bn - 1) / bn);
      f.name = "rn conv fwd (halo)"; f.N = u.kout; f.block_n = bn; f.b_mn = true; f.b_rows_per_tap = u.g.C; f.tma_out = 1;
      f.epi.out_bf16 = s.data[u.z];
      if ((rc = make_map_4d_box(&f.a, s.data[u.in], u.g.N, u.g.H, u.g.W, u.g.C, f.hs.PW, f.hs.TH + 2))) return rc;
      if ((rc = make_map(&f.b, kind, wlp(u.v_w), u.kout, rsc, u.kout, bn, true))) return rc;
      if ((rc = make_map_4d_box(&f.c, s.data[u.z], u.g.N, u.g.OH, u.g.OW, u.kout, f.hs.PW, f.hs.TH))) return rc;
      HaloConv& d = s.hdgrad[i];
      fill_halo_shape(&d.hs, u.g.N, u.g.H, u.g.W, u.kout, 1, (u.g.C + bnd - 1) / bnd);
      d.name = "rn conv dgrad (halo)"; d.N = u.g.C; d.block_n = bnd; d.b_mn = false; d.b_rows_per_tap = u.g.C;
      d.epi.out_bf16 = s.grad[u.in];
      if ((rc = make_map_4d_box(&d.a, s.grad[u.z], u.g.N, u.g.OH, u.g.OW, u.kout, d.hs.PW, d.hs.TH + 2))) return rc;
      if ((rc = make_map(&d.b, kind, wlp(u.v_w), rsc, u.kout, u.kout, bnd, false))) return rc;
      if ((rc = make_map_4d_box(&d.c, s.grad[u.in], u.g.N, u.g.H, u.g.W, u.g.C, d.hs.PW, d.hs.TH))) return rc;
    }
    if (u.implicit) {
      const int bn = u.kout >= 128 ? 128 : 64;
      int tn, th;
      // forward: tiles of output pixels, A = boxes of the input at each tap
      ConvGemm& f = s.ifwd[i];
      pick_conv_tile(u.g.N, u.g.OH, u.g.OW, false, &tn, &th);
      fill_conv_shape(&f.cs, u.g.N, u.g.OH, u.g.OW, tn, th, u.g.R, u.g.S, u.g.stride, u.g.pad, u.g.C, 0);
      f.name = "rn conv fwd (implicit)"; f.N = u.kout; f.ldc = u.kout; f.block_n = bn; f.b_mn = true;
      f.b_rows_per_tap = u.g.C; f.C = nullptr;
      f.epi = tc::Epilogue{nullptr, nullptr, 0, s.data[u.z], 1};
      if ((rc = make_map_4d(&f.a, s.data[u.in], u.g.N, u.g.H, u.g.W, u.g.C, u.g.OW, th, tn, u.g.stride))) return rc;
      if ((rc = make_map(&f.b, kind, wlp(u.v_w), u.kout, rsc, u.kout, bn, true))) return rc;
      if (tma_out_enabled()) {
        f.tma_out = 1;
        if ((rc = make_map_4d(&f.c, s.data[u.z], u.g.N, u.g.OH, u.g.OW, u.kout, u.g.OW, th, tn, 1))) return rc;
      }
      // weight gradient: pixel tiles are the K dimension
      ConvWgrad& w = s.iwgrad[i];
      pick_conv_tile(u.g.N, u.g.OH, u.g.OW, true, &tn, &th);
      fill_conv_shape(&w.cs, u.g.N, u.g.OH, u.g.OW, tn, th, u.g.R, u.g.S, u.g.stride, u.g.pad, u.g.C, 0);
      const long long tiles = ((rsc + 127) / 128) * ((u.kout + bn - 1) / bn);
      w.name = "rn conv wgrad (implicit)"; w.M = rsc; w.N = u.kout; w.block_n = bn; w.C = grd(u.v_w); w.partial = s.splitk;
      w.splits = (int)wgrad_splits(tiles);
      if ((rc = make_map_4d(&w.x, s.data[u.in], u.g.N, u.g.H, u.g.W, u.g.C, u.g.OW, th, tn, u.g.stride))) return rc;
      if ((rc = make_map_4d(&w.dz, s.grad[u.z], u.g.N, u.g.OH, u.g.OW, u.kout, u.g.OW, th, tn, 1))) return rc;
      // data gradient, stride 1: tiles of input pixels, A = boxes of dz at the mirrored tap offsets
      if (u.g.stride == 1) {
        ConvGemm& d = s.idgrad[i];
        const int bnd = u.g.C >= 128 ? 128 : 64;
        pick_conv_tile(u.g.N, u.g.H, u.g.W, false, &tn, &th);
        fill_conv_shape(&d.cs, u.g.N, u.g.H, u.g.W, tn, th, u.g.R, u.g.S, 1, u.g.pad, u.kout, 1);
        d.name = "rn conv dgrad (implicit)"; d.N = u.g.C; d.ldc = u.g.C; d.block_n = bnd; d.b_mn = false;
        d.b_rows_per_tap = u.g.C; d.C = nullptr;
        d.epi = tc::Epilogue{nullptr, nullptr, 0, nullptr, 1};   // out_bf16 chosen per call
        if ((rc = make_map_4d(&d.a, s.grad[u.z], u.g.N, u.g.OH, u.g.OW, u.kout, u.g.W, th, tn, 1))) return rc;
        if ((rc = make_map(&d.b, kind, wlp(u.v_w), rsc, u.kout, u.kout, bnd, false))) return rc;
        if (tma_out_enabled()) {   // store or reduce-add (chosen per call) straight into the input's gradient
          d.tma_out = 1;
          if ((rc = make_map_4d(&d.c, s.grad[u.in], u.g.N, u.g.H, u.g.W, u.g.C, u.g.W, th, tn, 1))) return rc;
        }
      }
      // data gradient, stride 2: one GEMM per parity class of the input position, tiles over the
      // OUTPUT pixel space (A = boxes of dz at the class's tap offsets), stored through a map of the
      // input gradient with element strides 2 -- no dcol, no col2im
      if (u.g.stride == 2 && u.in != m->t_image && tma_out_enabled() && s2_dgrad_enabled() && u.g.H == 2 * u.g.OH &&
          u.g.W == 2 * u.g.OW && 2 * u.g.OW <= 256) {
        const int bnd = u.g.C >= 128 ? 128 : 64;
        pick_conv_tile(u.g.N, u.g.OH, u.g.OW, false, &tn, &th);
        for (int cls = 0; cls < 4; ++cls) {
          ConvGemm& d = s.s2dgrad[i][cls];
          fill_conv_shape(&d.cs, u.g.N, u.g.OH, u.g.OW, tn, th, 1, 1, 1, 0, u.kout, 0);
          if (fill_dgrad_s2_class(&d.cs, u.g.R, cls >> 1, cls & 1) == 0) continue;
          d.name = "rn conv dgrad (stride 2)"; d.N = u.g.C; d.ldc = u.g.C; d.block_n = bnd; d.b_mn = false;
          d.b_rows_per_tap = u.g.C; d.C = nullptr; d.tma_out = 1;
          d.epi = tc::Epilogue{nullptr, nullptr, 0, s.grad[u.in], 1};
          if ((rc = make_map_4d(&d.a, s.grad[u.z], u.g.N, u.g.OH, u.g.OW, u.kout, u.g.OW, th, tn, 1))) return rc;
          if ((rc = make_map(&d.b, kind, wlp(u.v_w), rsc, u.kout, u.kout, bnd, false))) return rc;
          if ((rc = make_map_4d(&d.c, s.grad[u.in], u.g.N, u.g.H, u.g.W, u.g.C, u.g.OW, th, tn, 2))) return rc;
        }
      }
    }
    // explicit-operand descriptors: direct units, the stem, and the stride-2 data gradients
    const void* a_ptr = u.direct ? static_cast<const void*>(s.data[u.in])
                                 : (u.implicit ? static_cast<const void*>(s.dcol) : static_cast<const void*>(s.col[i]));
    const long long a_ld = u.direct ? u.g.C : u.g.Kp;
    const int bn = u.kout >= 128 ? 128 : 64;
    // forward: z = col x W
    TcGemm& f = s.fwd[i];
    f.name = "rn conv fwd"; f.M = (int)M; f.N = u.kout; f.K = rsc; f.C = nullptr; f.ldc = u.kout; f.block_n = bn;
    f.splits = 1; f.a_mn = false; f.b_mn = true;
    f.epi = tc::Epilogue{nullptr, nullptr, 0, s.data[u.z], 0};
    if ((rc = make_map(&f.a, kind, a_ptr, M, rsc, a_ld, tc::kBlockM, false))) return rc;
    if ((rc = make_map(&f.b, kind, wlp(u.v_w), u.kout, rsc, u.kout, bn, true))) return rc;
    if (tma_out_enabled()) {
      f.tma_out = 1;
      if ((rc = make_map(&f.c, kind, s.data[u.z], M, u.kout, u.kout, tc::kBlockM, false))) return rc;
    }
    // wgrad: dW = col^T x dz
    TcGemm& w = s.wgrad[i];
    const long long tiles = ((rsc + 127) / 128) * ((u.kout + bn - 1) / bn);
    w.name = "rn conv wgrad"; w.M = rsc; w.N = u.kout; w.K = (int)M; w.C = grd(u.v_w); w.ldc = u.kout; w.block_n = bn;
    w.splits = (int)wgrad_splits(tiles);
    w.partial = s.splitk; w.a_mn = true; w.b_mn = true;
    if ((rc = make_map(&w.a, kind, a_ptr, rsc, M, a_ld, tc::kBlockM, true))) return rc;
    if ((rc = make_map(&w.b, kind, s.grad[u.z], u.kout, M, u.kout, bn, true))) return rc;
    // dgrad: dcol = dz x W^T (not for the stem: the image needs no gradient)
    if (u.in != m->t_image) {
      TcGemm& d = s.dgrad[i];
      const int bnd = rsc >= 128 ? 128 : 64;
      d.name = "rn conv dgrad"; d.M = (int)M; d.N = rsc; d.K = u.kout; d.C = nullptr; d.block_n = bnd; d.splits = 1;
      d.a_mn = false; d.b_mn = false;
      d.ldc = u.direct ? u.g.C : u.g.Kp;
      d.epi = tc::Epilogue{nullptr, nullptr, 0, nullptr, 0};   // out_bf16 chosen per call (grad buffer or staging)
      if ((rc = make_map(&d.a, kind, s.grad[u.z], M, u.kout, u.kout, tc::kBlockM, false))) return rc;
      if ((rc = make_map(&d.b, kind, wlp(u.v_w), rsc, u.kout, u.kout, bnd, false))) return rc;
      if (tma_out_enabled()) {   // 1x1: into the input's gradient (store or reduce-add per call); else the dcol staging
        d.tma_out = 1;
        if ((rc = make_map(&d.c, kind, u.direct ? s.grad[u.in] : s.dcol, M, rsc, d.ldc, tc::kBlockM, false))) return rc;
      }
    }
  }
  const int B = m->B, NC = m->classes;
  TcGemm& ff = s.fc_fwd;
  ff.name = "rn fc fwd"; ff.M = B; ff.N = NC; ff.K = 2048; ff.C = s.logits; ff.ldc = NC; ff.block_n = 128; ff.splits = 1;
  ff.a_mn = false; ff.b_mn = true;
  ff.epi = tc::Epilogue{j->params + j->vars[m->v_fcb].offset, nullptr, 0, nullptr, 0};
  if ((rc = make_map(&ff.a, kind, s.pooled, B, 2048, 2048, tc::kBlockM, false))) return rc;
  if ((rc = make_map(&ff.b, kind, wlp(m->v_fcw), NC, 2048, NC, 128, true))) return rc;
  TcGemm& fw = s.fc_wgrad;
  fw.name = "rn fc wgrad"; fw.M = 2048; fw.N = NC; fw.K = B; fw.C = grd(m->v_fcw); fw.ldc = NC; fw.block_n = 128;
  fw.splits = 1; fw.a_mn = true; fw.b_mn = true;
  if ((rc = make_map(&fw.a, kind, s.pooled, 2048, B, 2048, tc::kBlockM, true))) return rc;
  if ((rc = make_map(&fw.b, kind, s.dlogits, NC, B, NC, 128, true))) return rc;
  TcGemm& fd = s.fc_dgrad;
  fd.name = "rn fc dgrad"; fd.M = B; fd.N = 2048; fd.K = NC; fd.C = nullptr; fd.ldc = 2048; fd.block_n = 128; fd.splits = 1;
  fd.a_mn = false; fd.b_mn = false;
  fd.epi = tc::Epilogue{nullptr, nullptr, 0, s.dpooled, 0};
  if ((rc = make_map(&fd.a, kind, s.dlogits, B, NC, NC, tc::kBlockM, false))) return rc;
  if ((rc = make_map(&fd.b, kind, wlp(m->v_fcw), 2048, NC, NC, 128, false))) return rc;
  s.maps_ready = true;
  return TFG_OK;
}

static int run_im2col(tfg_job* j, const RnUnit& u, const bf16* in, bf16* col) {
  const long long M = (long long)u.g.N * u.g.OH * u.g.OW;
  const int rsc = u.g.R * u.g.S * u.g.C;
  if (u.g.C % 8 == 0) {
    rn::im2col_kernel<8><<<grid_for(M * u.g.R * u.g.S * (u.g.C / 8)), 256, 0, j->stream>>>(in, col, u.g);
  } else {
    rn::im2col_rows_kernel<<<grid_for(M * (u.g.Kp / 8)), 256, 0, j->stream>>>(in, col, u.g);
    RN_OK(j, "rn im2col (thin)");
    return TFG_OK;
  }
  RN_OK(j, "rn im2col");
  if (u.g.Kp != rsc) {
    rn::zero_pad_cols_kernel<<<grid_for(M * (u.g.Kp - rsc)), 256, 0, j->stream>>>(col, M, rsc, u.g.Kp);
    RN_OK(j, "rn im2col pad");
  }
  return TFG_OK;
}

static void bn_chunks(long long M, int C, int* chunks, int* rows) {
  const int groups = C / 64;
  int c = (kBnTargetBlocks + groups - 1) / groups;
  c = (int)std::max<long long>(1, std::min<long long>(c, (M + 31) / 32));
  int r = (int)((M + c - 1) / c);
  *chunks = (int)((M + r - 1) / r);
  *rows = r;
}

// batch-norm statistics come from the GEMM epilogue's column sums wherever the output went through the TMA
// path (measured against a pass over z: profiles/README.md, round 1)
static constexpr bool bn_fused_enabled() { return true; }

// batch statistics of unit i's convolution output: folded from the GEMM epilogue's per-tile column
// sums when the output went through the TMA path, else reduced from z
static int bn_forward_stats(tfg_job* j, RnSlot& s, const RnUnit& u, int i, bool fused, bool update_moving) {
  const long long M = (long long)u.g.N * u.g.OH * u.g.OW;
  float* mm = update_moving ? u.mov_mean : nullptr;
  float* mv = update_moving ? u.mov_var : nullptr;
  const int groups = u.kout / 64;
  if (fused) {
    rn::bn_fold_kernel<<<groups, 1024, 0, j->stream>>>(s.tile_stats, stat_tiles(u), M, u.kout, 1e-5f, 0.997f, s.mean[i],
                                                       s.invstd[i], mm, mv);
    RN_OK(j, "rn bn fold");
    return TFG_OK;
  }
  int chunks, rows;
  bn_chunks(M, u.kout, &chunks, &rows);
  rn::bn_stats_kernel<<<dim3(groups, chunks), 256, 0, j->stream>>>(s.data[u.z], s.bn_partial, s.bn_done, M, u.kout, rows,
                                                                  1e-5f, 0.997f, s.mean[i], s.invstd[i], mm, mv);
  RN_OK(j, "rn bn stats");
  return TFG_OK;
}

int resnet_forward_backward(tfg_job* j, int slot, const float* images, const int32_t* labels) {
  ResnetModel* m = j->resnet;
  RnSlot& s = m->slots[slot];
  cudaStream_t st = j->stream;
  const float* P = j->params;
  float* G = j->grads[slot];
  auto par = [&](int v) { return P + j->vars[v].offset; };
  auto grd = [&](int v) { return G + j->vars[v].offset; };
  int rc = prepare_maps(j, slot);
  if (rc) return rc;
  const int B = m->B;
  const bool update_moving = slot == 0;

  // ================= forward =================
  nvtxRangePushA("resnet50 forward");
  {
    const long long n = m->tensors[m->t_image].elems();
    rn::to_bf16_kernel<<<grid_for(n), 256, 0, st>>>(images, s.data[m->t_image], n);
    RN_OK(j, "rn image to bf16");
  }
  for (size_t i = 0; i < m->units.size(); ++i) {
    const RnUnit& u = m->units[i];
    const long long M = (long long)u.g.N * u.g.OH * u.g.OW;
    if (u.halo) {
      HaloConv f = s.hfwd[i];
      const bool fused = bn_fused_enabled();
      if (fused) f.epi.col_stats = s.tile_stats;
      if ((rc = launch_conv_halo(j, f, st))) return rc;
      if ((rc = bn_forward_stats(j, s, u, (int)i, fused, update_moving))) return rc;
    } else if (u.implicit) {
      ConvGemm f = s.ifwd[i];
      const bool fused = f.tma_out != 0 && bn_fused_enabled();
      if (fused) f.epi.col_stats = s.tile_stats;
      if ((rc = launch_conv_gemm(j, f, st))) return rc;
      if ((rc = bn_forward_stats(j, s, u, (int)i, fused, update_moving))) return rc;
    } else {
      if (!u.direct && (rc = run_im2col(j, u, s.data[u.in], s.col[i]))) return rc;
      TcGemm f = s.fwd[i];
      const bool fused = f.tma_out != 0 && bn_fused_enabled();
      if (fused) f.epi.col_stats = s.tile_stats;
      if ((rc = launch_tc(j, f, TFG_MATH_BF16, st))) return rc;
      if ((rc = bn_forward_stats(j, s, u, (int)i, fused, update_moving))) return rc;
    }
    rn::bn_apply_kernel<<<grid_resident(M * u.kout / 8), 256, 0, st>>>(
        s.data[u.z], s.mean[i], s.invstd[i], par(u.v_gamma), par(u.v_beta), u.res >= 0 ? s.data[u.res] : nullptr,
        s.data[u.y], u.relu ? s.relu_bits[i] : nullptr, M * u.kout, u.kout);
    RN_OK(j, "rn bn apply");
    if (u.y == m->t_pool_in) {
      const RnTensor& ti = m->tensors[m->t_pool_in];
      const RnTensor& to = m->tensors[m->t_pool_out];
      rn::maxpool_fwd_kernel<<<grid_for(to.elems() / 8), 256, 0, st>>>(s.data[m->t_pool_in], s.data[m->t_pool_out], s.pool_arg,
                                                                  (int)ti.n, (int)ti.h, (int)ti.w, (int)ti.c, (int)to.h,
                                                                  (int)to.w);
      RN_OK(j, "rn maxpool fwd");
    }
  }
  const RnTensor& tl = m->tensors[m->t_last];
  const int hw = (int)(tl.h * tl.w);
  rn::avgpool_fwd_kernel<<<(B * 2048 + 255) / 256, 256, 0, st>>>(s.data[m->t_last], s.pooled, B, hw, 2048);
  RN_OK(j, "rn avgpool fwd");
  if ((rc = launch_tc(j, s.fc_fwd, TFG_MATH_BF16, st))) return rc;
  rn::softmax_xent_wide_kernel<<<(B + 7) / 8, 256, 0, st>>>(s.logits, labels, s.dlogits, s.loss_vec, B, m->classes);
  RN_OK(j, "rn softmax xent");
  rn::mean_kernel<<<1, 256, 0, st>>>(s.loss_vec, j->loss + j->local_replicas[slot], B);
  RN_OK(j, "rn loss mean");

  nvtxRangePop();
  // ================= backward =================
  NvtxRange nvtx_bwd("resnet50 backward");
  std::vector<char> written(m->tensors.size(), 0);   // has this tensor's gradient buffer been written yet?
  if ((rc = launch_tc(j, s.fc_wgrad, TFG_MATH_BF16, st))) return rc;
  rn::colsum_bf16_kernel<<<(m->classes + 255) / 256, 256, 0, st>>>(s.dlogits, grd(m->v_fcb), B, m->classes);
  RN_OK(j, "rn fc bias grad");
  if ((rc = launch_tc(j, s.fc_dgrad, TFG_MATH_BF16, st))) return rc;
  rn::avgpool_bwd_kernel<<<grid_for(tl.elems()), 256, 0, st>>>(s.dpooled, s.grad[m->t_last], B, hw, 2048);
  RN_OK(j, "rn avgpool bwd");
  written[m->t_last] = 1;

  for (int i = (int)m->units.size() - 1; i >= 0; --i) {
    const RnUnit& u = m->units[i];
    const long long M = (long long)u.g.N * u.g.OH * u.g.OW;
    if (u.y == m->t_pool_in) {   // gradient arrives through the max pool
      const RnTensor& ti = m->tensors[m->t_pool_in];
      const RnTensor& to = m->tensors[m->t_pool_out];
      rn::maxpool_bwd_kernel<<<grid_for(ti.elems() / 8), 256, 0, st>>>(s.grad[m->t_pool_out], s.pool_arg, s.grad[m->t_pool_in],
                                                                  (int)ti.n, (int)ti.h, (int)ti.w, (int)ti.c, (int)to.h,
                                                                  (int)to.w);
      RN_OK(j, "rn maxpool bwd");
      written[m->t_pool_in] = 1;
    }
    if (!written[u.y]) return fail(j, TFG_ESTATE, "internal: gradient of " + u.name + "/y was never produced");
    int chunks, rows;
    bn_chunks(M, u.kout, &chunks, &rows);
    bf16* g_out = nullptr;
    int g_acc = 0;
    if (u.res >= 0) { g_out = s.grad[u.res]; g_acc = written[u.res]; written[u.res] = 1; }
    const int agrid = grid_resident(M * u.kout / 8);
    if (u.relu)
      rn::bn_bwd_stats_kernel<true><<<agrid, 256, 0, st>>>(s.grad[u.y], s.relu_bits[i], s.data[u.z], s.mean[i], s.tile_stats,
                                                           M * u.kout, u.kout);
    else
      rn::bn_bwd_stats_kernel<false><<<agrid, 256, 0, st>>>(s.grad[u.y], nullptr, s.data[u.z], s.mean[i], s.tile_stats,
                                                            M * u.kout, u.kout);
    RN_OK(j, "rn bn bwd stats");
    rn::bn_bwd_fold_kernel<<<u.kout / 64, 1024, 0, st>>>(s.tile_stats, agrid, u.kout, s.invstd[i], grd(u.v_beta),
                                                         grd(u.v_gamma));
    RN_OK(j, "rn bn bwd fold");
    if (u.relu)
      rn::bn_bwd_apply_kernel<true><<<agrid, 256, 0, st>>>(s.grad[u.y], s.relu_bits[i], s.data[u.z], s.mean[i], s.invstd[i],
                                                           par(u.v_gamma), grd(u.v_beta), grd(u.v_gamma), s.grad[u.z], g_out,
                                                           g_acc, M * u.kout, u.kout, M);
    else
      rn::bn_bwd_apply_kernel<false><<<agrid, 256, 0, st>>>(s.grad[u.y], nullptr, s.data[u.z], s.mean[i], s.invstd[i],
                                                            par(u.v_gamma), grd(u.v_beta), grd(u.v_gamma), s.grad[u.z], g_out,
                                                            g_acc, M * u.kout, u.kout, M);
    RN_OK(j, "rn bn bwd apply");
    // weight gradient: a leaf of the dependency graph (nothing downstream in this step reads it before
    // the apply), so it goes to the side stream and overlaps the data-gradient / batch-norm chain; the
    // side stream runs the weight gradients one after another, which keeps their shared split-K
    // partial buffer safe.  (the im2col operand of the stem is the one the forward pass built)
    {
      cudaStream_t ws = side_fork(j);
      if (u.implicit) {
        if ((rc = launch_conv_wgrad(j, s.iwgrad[i], ws))) return rc;
      } else if ((rc = launch_tc(j, s.wgrad[i], TFG_MATH_BF16, ws))) {
        return rc;
      }
    }
    if (u.in == m->t_image) continue;
    // data gradient
    TcGemm d = s.dgrad[i];
    const long long in_elems = m->tensors[u.in].elems();
    if (u.halo) {
      HaloConv dg = s.hdgrad[i];
      dg.tma_out = written[u.in] ? 2 : 1;
      if ((rc = launch_conv_halo(j, dg, st))) return rc;
    } else if (u.implicit && u.g.stride == 1) {
      ConvGemm dg = s.idgrad[i];
      if (dg.tma_out) {
        dg.epi.out_bf16 = s.grad[u.in];
        dg.tma_out = written[u.in] ? 2 : 1;
        if ((rc = launch_conv_gemm(j, dg, st))) return rc;
      } else {
        dg.epi.out_bf16 = written[u.in] ? s.dcol : s.grad[u.in];
        if ((rc = launch_conv_gemm(j, dg, st))) return rc;
        if (written[u.in]) {
          rn::add_bf16_kernel<<<grid_for(in_elems / 8), 256, 0, st>>>(s.dcol, s.grad[u.in], in_elems, 1);
          RN_OK(j, "rn grad add");
        }
      }
    } else if (u.direct) {
      if (d.tma_out) {
        d.epi.out_bf16 = s.grad[u.in];
        d.tma_out = written[u.in] ? 2 : 1;
        if ((rc = launch_tc(j, d, TFG_MATH_BF16, st))) return rc;
      } else if (!written[u.in]) {
        d.epi.out_bf16 = s.grad[u.in];
        if ((rc = launch_tc(j, d, TFG_MATH_BF16, st))) return rc;
      } else {
        d.epi.out_bf16 = s.dcol;
        if ((rc = launch_tc(j, d, TFG_MATH_BF16, st))) return rc;
        rn::add_bf16_kernel<<<grid_for(in_elems / 8), 256, 0, st>>>(s.dcol, s.grad[u.in], in_elems, 1);
        RN_OK(j, "rn grad add");
      }
    } else if (u.implicit && (s.s2dgrad[i][0].cs.ntaps > 0)) {
      // stride 2: the parity classes write disjoint positions; a 1x1 convolution reaches the
      // (even, even) class only, so a gradient buffer nobody wrote yet is cleared first
      const bool partial_cover = s.s2dgrad[i][3].cs.ntaps == 0;
      if (!written[u.in] && partial_cover) {
        cudaError_t e = cudaMemsetAsync(s.grad[u.in], 0, sizeof(bf16) * (size_t)in_elems, st);
        if (e != cudaSuccess) return fail_cuda(j, e, "cudaMemsetAsync(input gradient)");
      }
      for (int cls = 0; cls < 4; ++cls) {
        ConvGemm dg = s.s2dgrad[i][cls];
        if (dg.cs.ntaps == 0) continue;
        dg.tma_out = (written[u.in] || partial_cover) ? 2 : 1;
        if ((rc = launch_conv_gemm(j, dg, st))) return rc;
      }
    } else {
      d.epi.out_bf16 = s.dcol;
      if ((rc = launch_tc(j, d, TFG_MATH_BF16, st))) return rc;
      if (u.g.C % 8 == 0)
        rn::col2im_kernel<8><<<grid_for(in_elems / 8), 256, 0, st>>>(s.dcol, s.grad[u.in], u.g, written[u.in]);
      else
        rn::col2im_kernel<1><<<grid_for(in_elems), 256, 0, st>>>(s.dcol, s.grad[u.in], u.g, written[u.in]);
      RN_OK(j, "rn col2im");
    }
    written[u.in] = 1;
  }
  side_join(j);
  return TFG_OK;
}

int resnet_logits(tfg_job* j, int slot, float* host, size_t bytes) {
  ResnetModel* m = j->resnet;
  if (bytes != sizeof(float) * (size_t)m->B * m->classes) return fail(j, TFG_EINVAL, "logits byte count mismatch");
  TFG_CUDA(j, cudaMemcpyAsync(host, m->slots[slot].logits, bytes, cudaMemcpyDeviceToHost, j->stream));
  TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  return TFG_OK;
}

// debug / parity access to an activation ("<unit>/z", "<unit>/y", "pool1", "image") or, with a
// "grad:" prefix, to its gradient; converted to fp32 on the host
int resnet_tensor_get(tfg_job* j, int slot, const char* name, float* host, size_t cap, size_t* n_out) {
  ResnetModel* m = j->resnet;
  std::string nm = name;
  bool want_grad = false;
  if (nm.rfind("grad:", 0) == 0) { want_grad = true; nm = nm.substr(5); }
  for (size_t t = 0; t < m->tensors.size(); ++t) {
    if (m->tensors[t].name != nm) continue;
    const size_t n = (size_t)m->tensors[t].elems();
    if (n_out) *n_out = n;
    if (cap < n) return fail(j, TFG_ELIMIT, "host buffer too small for tensor " + nm);
    const bf16* src = want_grad ? m->slots[slot].grad[t] : m->slots[slot].data[t];
    if (!src) return fail(j, TFG_EINVAL, "tensor " + nm + " has no such buffer");
    std::vector<uint16_t> tmp(n);
    TFG_CUDA(j, cudaMemcpyAsync(tmp.data(), src, n * 2, cudaMemcpyDeviceToHost, j->stream));
    TFG_CUDA(j, cudaStreamSynchronize(j->stream));
    for (size_t i = 0; i < n; ++i) {
      uint32_t bits = (uint32_t)tmp[i] << 16;
      memcpy(&host[i], &bits, 4);
    }
    return TFG_OK;
  }
  return fail(j, TFG_EINVAL, "no tensor named " + nm);
}

}  // namespace tfg
