// Packed parameter layout and workspace carving for PoseNet / PoseRefineNet (host side).
#pragma once
#include "common.cuh"

namespace plr2 {

inline size_t al32(size_t n) { return (n + 31) / 32 * 32; }   // 128-byte granularity
inline int al4(int n) { return (n + 3) / 4 * 4; }

// Packed weights (floats).  Heads are concatenated so one launch covers r/t/c:
//   PoseNet : h1pf [1920,384] = conv1_{r,t,c}.weight[:, 0:384]   (per-point part of the 1408 inputs)
//             h1ap [1920,1024] = conv1_{r,t,c}.weight[:, 384:1408] (per-frame part: global feature)
//             h2w [3][256,640], h3w [3][128,256], h4w [8*O,128] (r rows, t rows, c rows)
//   Refiner : h1ap [1024,1024] = conv1_{r,t}.weight, h2w [2][128,512], h4w [7*O,128]; no h1pf/h3w
struct NetLayout {
  int net, O, heads;          // heads = 3 (posenet) / 2 (refiner)
  int c5k;                    // conv5 input channels: 256 / 384
  int h1n, h2n, h2k, h3n, h3k, h4k;
  int oc[3], on[3], ldo;      // column offset / width of each head in the raw output buffer
  size_t c1w, e1w, c2w, e2w, c5w, c6w, h1pf, h1ap, h2w, h3w, h4w;
  size_t c1b, e1b, c2b, e2b, c5b, c6b, h1b, h2b, h3b, h4b;
  size_t total;
};

inline NetLayout make_layout(int net, int O) {
  NetLayout L{};
  L.net = net; L.O = O;
  const bool pn = net == PLR2_NET_POSENET;
  L.heads = pn ? 3 : 2;
  L.c5k = pn ? 256 : 384;
  L.h1n = pn ? 640 : 512;          // per head
  L.h2n = pn ? 256 : 128; L.h2k = L.h1n;
  L.h3n = pn ? 128 : 0;   L.h3k = L.h2n;
  L.h4k = 128;
  const int widths[3] = {4, 3, 1};
  int c = 0;
  for (int h = 0; h < 3; ++h) {
    L.oc[h] = c; L.on[h] = h < L.heads ? widths[h] * O : 0;
    c += al4(L.on[h]);
  }
  L.ldo = c;
  size_t o = 0;
  auto take = [&](size_t n) { size_t r = o; o += al32(n); return r; };
  L.c1w = take(64 * 4); L.e1w = take(64 * 32); L.c2w = take(128 * 64); L.e2w = take(128 * 64);
  L.c5w = take((size_t)512 * L.c5k); L.c6w = take(1024 * 512);
  L.h1pf = take(pn ? (size_t)3 * 640 * 384 : 0);
  L.h1ap = take((size_t)L.heads * L.h1n * 1024);
  L.h2w = take((size_t)L.heads * L.h2n * L.h2k);
  L.h3w = take((size_t)L.heads * L.h3n * L.h3k);
  L.h4w = take((size_t)(L.on[0] + L.on[1] + L.on[2]) * 128);
  L.c1b = take(64); L.e1b = take(64); L.c2b = take(128); L.e2b = take(128); L.c5b = take(512); L.c6b = take(1024);
  L.h1b = take((size_t)L.heads * L.h1n); L.h2b = take((size_t)L.heads * L.h2n); L.h3b = take((size_t)L.heads * L.h3n);
  L.h4b = take((size_t)L.on[0] + L.on[1] + L.on[2]);
  L.total = o;
  return L;
}

// Workspace (floats) for R = frames*n_points rows.
struct NetWorkspace {
  unsigned int *status;    // first 128 bytes of every network workspace: input-range flags of the last forward (PLR2_STATUS_*)
  float *pf, *x5, *x6, *ap, *hb, *h1, *h2, *h3, *out, *embT, *dpf, *dap, *dhb;
  size_t red_cap = 0;      // floats in `red`
  float *red;              // training: [BGRAD_SLICES][BGRAD_MAX_COLS] partial column sums of the bias gradients
  // training: scratch of the tensor-core contractions (gemm_tc_f32.cu): split-bf16 planes of the two operands
  // (tc_pa, tc_pb: 2 planes x tc_elems bf16 each) and an FP32 output staging area (tc_ct: tc_elems floats)
  float *tc_pa, *tc_pb, *tc_ct;
  size_t tc_elems;
  size_t floats;
};

constexpr int BGRAD_SLICES = 64, BGRAD_MAX_COLS = 2048;

inline NetWorkspace carve_workspace(const NetLayout &L, void *base, int frames, int n_points, int training) {
  NetWorkspace W{};
  const size_t R = (size_t)frames * n_points, F = frames;
  const bool pn = L.net == PLR2_NET_POSENET;
  float *p = reinterpret_cast<float *>(base);
  size_t o = 0;
  auto take = [&](size_t n) { float *r = p ? p + o : nullptr; o += al32(n); return r; };
  W.status = reinterpret_cast<unsigned int *>(take(32));
  W.pf = take(R * 384); W.x5 = take(R * 512); W.x6 = take(R * 1024);
  W.ap = take(F * 1024);
  const size_t h1all = (size_t)L.heads * L.h1n, h2all = (size_t)L.heads * L.h2n, h3all = (size_t)L.heads * L.h3n;
  if (pn) {
    W.hb = take(F * h1all); W.h1 = take(R * h1all); W.h2 = take(R * h2all); W.h3 = take(R * h3all);
    W.out = take(R * L.ldo);
  } else {
    W.hb = nullptr; W.h1 = take(F * h1all); W.h2 = take(F * h2all); W.h3 = nullptr;
    W.out = take(F * L.ldo);
  }
  W.embT = take(R * 32);
  if (training) {
    W.dpf = take(R * 384); W.dap = take(F * 1024); W.dhb = take(pn ? F * h1all : 0);
    // (or one partial per 32-row tile when the tensor-core split of dY produces them: gemm.cuh colsum_part)
    W.red_cap = (size_t)(BGRAD_SLICES > (R + 31) / 32 ? BGRAD_SLICES : (R + 31) / 32) * BGRAD_MAX_COLS;
    W.red = take(W.red_cap);
    // largest operand / output of any contraction: 1920 columns (conv1_{r,t,c}) x rows padded to the K tile,
    // or the 1920 x 1024 per-frame weight block
    const size_t rp = (R + 63) / 64 * 64;
    W.tc_elems = (size_t)1920 * ((rp > 1024 ? rp : 1024) + 2048);   // + K padding of up to 32 split-K chunks
    W.tc_pa = take(W.tc_elems); W.tc_pb = take(W.tc_elems); W.tc_ct = take(W.tc_elems);   // 2 bf16 planes = tc_elems floats
  }
  W.floats = o;
  return W;
}


// Workspace of the tensor-core (split-bf16) inference path.  Activations consumed by tcgen05 layers
// are stored as two bf16 planes [2][R][width]; *_plane is the plane stride in elements.
struct TcWorkspace {
  unsigned int *status;                    // same place as NetWorkspace::status
  void *pf, *x5, *h1, *h2;                 // bf16 planes
  long long pf_plane, x5_plane, h1_plane, h2_plane;
  float *pool, *ap, *hb, *h3, *out;        // fp32; refiner: hb = l1 out, h3 = l2 out
  void *apb; long long apb_plane;          // pooled features as split-bf16 planes [2][F][1024] (operand of the M = frames GEMMs)
  size_t bytes;
};

inline TcWorkspace carve_tc_workspace(const NetLayout &L, void *base, int frames, int n_points) {
  TcWorkspace W{};
  const size_t R = (size_t)frames * n_points, F = frames;
  const bool pn = L.net == PLR2_NET_POSENET;
  uint8_t *p = reinterpret_cast<uint8_t *>(base);
  size_t o = 0;
  auto take = [&](size_t nbytes) { void *r = p ? p + o : nullptr; o += (nbytes + 127) / 128 * 128; return r; };
  const size_t h1all = (size_t)L.heads * L.h1n, h2all = (size_t)L.heads * L.h2n, h3all = (size_t)L.heads * L.h3n;
  W.status = reinterpret_cast<unsigned int *>(take(128));
  W.pf = take(2 * R * 384 * 2); W.pf_plane = (long long)R * 384;
  W.x5 = take(R * 512 * 2); W.x5_plane = 0;          // hi plane only: conv6 reads it through the pool (hi-only operand)
  W.pool = (float *)take(((R + 31) / 32) * 2 * 1024 * sizeof(float));
  W.ap = (float *)take(F * 1024 * sizeof(float));
  W.apb = take(2 * F * 1024 * 2); W.apb_plane = (long long)F * 1024;
  if (pn) {
    W.hb = (float *)take(F * h1all * sizeof(float));
    W.h1 = take(2 * R * h1all * 2); W.h1_plane = (long long)R * h1all;
    W.h2 = take(2 * R * h2all * 2); W.h2_plane = (long long)R * h2all;
    W.h3 = nullptr; W.out = nullptr;                  // conv3 output and the all-class head outputs are never materialised
    (void)h3all;
  } else {
    W.hb = (float *)take(F * h1all * sizeof(float));
    W.h3 = (float *)take(F * h2all * sizeof(float));
    W.out = (float *)take(F * L.ldo * sizeof(float));
  }
  W.bytes = o;
  return W;
}

}  // namespace plr2
