// Animation -> keypoints: the shield's human input without MuJoCo in the loop.
//
// Reference: `_control_human` (HRG/environments/manipulation/human_env.py:1710-1767) picks the frame of the env's current
// clip and writes the Pelvis mocap pose and 69 hinge angles into MuJoCo; `sim.forward()` runs the kinematic tree of
// HRG/models/assets/human/human.xml and `human_measurement` (human_env.py:459-463) reads the 23 site positions.
// Every body of that model has pos 0 / quat identity and three hinges z, y, x (XML order) about its own site, so with
// c_b the site of body b in the common rest frame, R_b its world rotation and a_b its site in the world:
//     a_b = a_p + R_p (c_b - c_p)          (the site is the joint anchor: it moves with the PARENT only)
//     R_b = R_p Rz(z_b) Ry(y_b) Rx(x_b)    (MuJoCo composes the joints of a body in XML order)
// with (R, a, c) of the Pelvis = (mocap rotation, mocap position, 0).
//
// with (R, a, c) of the Pelvis = (mocap rotation, mocap position, 0), and (human_env.py:1737-1763)
//     mocap position = Rc (p_t + offset) + human_pos_offset,  mocap rotation = Rp Rc Ra_t
// (Rc = human_base_quat * clip orientation, Rp = the env's human_rot_offset, p_t / Ra_t = the frame's Pelvis pose).
//
// B200 design: only Rp and human_pos_offset depend on the ENV; everything else depends on (clip, frame) alone. So the
// tree walk - 54 sincos and 18 matrix products per pose, FP64-pipe bound - is hoisted out of the tick: k_human_bake runs
// it ONCE per frame of the clip bank when the bank is created (lane = frame, parent frames in registers, frames that a
// later sibling subtree needs again - Pelvis and Chest - parked in one of kHumanSlots register slots), storing per frame
//     u = Rc (p_t + offset)   and   w_b = (Rc Ra_t) x (site b walked from a root at the origin).
// The per-tick kernel k_human_keypoints is then the gather + rigid transform  kp_b = (human_pos_offset + u) + Rp w_b:
// a block owns 32 consecutive envs, 32 lanes pick the frames (`_control_human`'s clock arithmetic, replay state update)
// and stage (pos + u, Rp) in shared memory, then all threads stream the 32 x 3K outputs - baked rows are read and the
// caller's [N,K,3] block is written as contiguous, fully coalesced runs. The bank is small (the reference's whole CMU
// set bakes to 11 MB) and stays in L2. HBM-bound: 24K + 24 row bytes in, 24K out, 96 placement, 24 clock/state per env.
// Measured history at 65 536 envs: walk per env per tick through shared-memory tiles 84.5 us; walk per env per tick
// from registers 35.7 us (profiles/r01f_ncu_summary.md); baked bank: see profiles/.
#pragma once
#include "state.cuh"

namespace hrg {

constexpr int kHumanBlock = 128;

constexpr int kHumanSlots = 3;       // parked frames; human.xml needs 2 (Pelvis, Chest)

struct HumanTreeDev {                // by value (constant bank; every index below is warp-uniform)
  int32_t K;
  int32_t order[HRG_MAX_KEYPOINTS];      // body visited i-th (depth-first preorder)
  int32_t has_child[HRG_MAX_KEYPOINTS];  // leaves need no rotation
  int32_t src[HRG_MAX_KEYPOINTS];        // -1: the parent's frame is the one in registers; s >= 0: fetch it from slot s
  int32_t save[HRG_MAX_KEYPOINTS];       // s >= 0: park this body's frame in slot s (a later sibling subtree needs it)
  double rel[HRG_MAX_KEYPOINTS][3];      // c_b - c_parent
};

constexpr int kHumanTileEnvs = 32;    // envs per block of k_human_keypoints

struct HumanDev {
  int32_t N, K, C, M;            // envs, bodies, channels per raw frame (7 + 3K), clips per env list
  int32_t n_clips;
  double step_length;
  const double* baked;           // [F][3 + 3K]: u, then w_b per body (k_human_bake)
  const int64_t* clip_off;       // [n_clips] first row
  const int32_t* clip_len;       // [n_clips]
  const double* clip_tab;        // [n_clips][12]: position_offset, then Rc = R(human_base_quat * orientation_quat) row-major
  double* placement;             // [N][12]: human_pos_offset, then Rp row-major
  int32_t* clip_list;            // [N][M]
  int32_t* list_idx;             // [N]  _human_animation_ids_index
  int64_t* anim_start;           // [N]  animation_start_time
};


// Host side: depth-first preorder of the tree (children in index order), leaf flags, anchor differences, and the
// register-slot plan. Returns 0, or -1 parents not before children, -2 more parked frames than kHumanSlots, -3 not
// connected to the Pelvis.
inline int human_build_tree(int K, const int32_t* parent, const double* anchor, HumanTreeDev* T) {
  for (int b = 0; b < K; ++b)
    if (parent[b] < -1 || parent[b] >= b) return -1;
  T->K = K;
  int stack[HRG_MAX_KEYPOINTS], top = 0, n = 0;
  for (int b = K - 1; b >= 0; --b)
    if (parent[b] < 0) stack[top++] = b;
  while (top > 0) {
    const int b = stack[--top];
    const int p = parent[b];
    T->order[n] = b;
    int kids = 0;
    for (int c = K - 1; c > b; --c)
      if (parent[c] == b) {
        stack[top++] = c;
        ++kids;
      }
    T->has_child[n] = kids > 0;
    for (int k = 0; k < 3; ++k) T->rel[n][k] = anchor[3 * b + k] - (p < 0 ? 0.0 : anchor[3 * p + k]);
    ++n;
  }
  if (n != K) return -3;
  // which frames must be parked: a body whose child is visited when the registers hold another body's frame
  bool parked[HRG_MAX_KEYPOINTS + 1] = {};  // index b + 1; 0 = the Pelvis
  int in_regs = -1;
  for (int i = 0; i < K; ++i) {
    const int b = T->order[i], p = parent[b];
    if (p != in_regs) parked[p + 1] = true;
    if (T->has_child[i]) in_regs = b;
  }
  // slot = number of parked proper ancestors (sibling subtrees are disjoint in time, nested ones are not)
  int slot_of[HRG_MAX_KEYPOINTS + 1];
  slot_of[0] = 0;  // the Pelvis always sits in slot 0
  for (int b = 0; b < K; ++b) {
    int s = 1;
    for (int a = parent[b]; a >= 0; a = parent[a]) s += parked[a + 1] ? 1 : 0;
    slot_of[b + 1] = s;
    if (parked[b + 1] && s >= kHumanSlots) return -2;
  }
  in_regs = -1;
  for (int i = 0; i < K; ++i) {
    const int b = T->order[i], p = parent[b];
    T->src[i] = p == in_regs ? -1 : slot_of[p + 1];
    T->save[i] = T->has_child[i] && parked[b + 1] ? slot_of[b + 1] : -1;
    if (T->has_child[i]) in_regs = b;
  }
  return 0;
}

HRG_HD void quat_xyzw_to_mat(double x, double y, double z, double w, double R[9]) {
  const double n = sqrt(x * x + y * y + z * z + w * w);  // scipy's from_quat / mj_kinematics normalise
  x /= n; y /= n; z /= n; w /= n;
  R[0] = 1.0 - 2.0 * (y * y + z * z); R[1] = 2.0 * (x * y - z * w);       R[2] = 2.0 * (x * z + y * w);
  R[3] = 2.0 * (x * y + z * w);       R[4] = 1.0 - 2.0 * (x * x + z * z); R[5] = 2.0 * (y * z - x * w);
  R[6] = 2.0 * (x * z - y * w);       R[7] = 2.0 * (y * z + x * w);       R[8] = 1.0 - 2.0 * (x * x + y * y);
}

// Per-clip constant of the root pose: position_offset, then R(human_base_quat * orientation_quat) - human_env.py:373,1750.
inline void human_clip_table(const double info[7], double tab[12]) {
  const double a[4] = {0.5, 0.5, 0.5, 0.5};  // human_base_quat (x,y,z,w)
  const double* b = info + 3;
  double q[4];                               // Hamilton product a * b, (x,y,z,w)
  q[3] = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
  q[0] = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
  q[1] = a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0];
  q[2] = a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3];
  for (int k = 0; k < 3; ++k) tab[k] = info[k];
  quat_xyzw_to_mat(q[0], q[1], q[2], q[3], tab + 3);
}

// 3x3 product with fused multiply-adds spelled out (the library is built with -fmad=false for the shield's parity-critical
// arithmetic; this path has a 1e-5 m tolerance and no exact-equality branches, so it takes the cheaper instruction).
HRG_HD void mat3_mul(const double A[9], const double B[9], double Cm[9]) {
  HRG_UNROLL
  for (int r = 0; r < 3; ++r)
    HRG_UNROLL
    for (int c = 0; c < 3; ++c) Cm[3 * r + c] = fma(A[3 * r], B[c], fma(A[3 * r + 1], B[3 + c], A[3 * r + 2] * B[6 + c]));
}

// Rz(z) Ry(y) Rx(x)
HRG_HD void hinge_zyx(double x, double y, double z, double R[9]) {
  double sx, cx, sy, cy, sz, cz;
  sincos(x, &sx, &cx); sincos(y, &sy, &cy); sincos(z, &sz, &cz);
  R[0] = cz * cy; R[1] = cz * sy * sx - sz * cx; R[2] = cz * sy * cx + sz * sx;
  R[3] = sz * cy; R[4] = sz * sy * sx + cz * cx; R[5] = sz * sy * cx - cz * sx;
  R[6] = -sy;     R[7] = cy * sx;                R[8] = cy * cx;
}

// The tree walk of one env: `row` = its frame row, `kp` = its 3K outputs (any stride-1 memory).
HRG_HD void human_tree_walk(const HumanTreeDev& T, const double* __restrict__ row, const double root[12],
                            double* __restrict__ kp) {
  double cur[12], slot[kHumanSlots][12];
  HRG_UNROLL
  for (int i = 0; i < 12; ++i) cur[i] = slot[0][i] = root[i];
  HRG_UNROLL
  for (int s = 1; s < kHumanSlots; ++s)
    HRG_UNROLL
    for (int i = 0; i < 12; ++i) slot[s][i] = 0.0;
  // the angles of the next body are fetched one visit ahead of the arithmetic that needs them
  int b = T.order[0];
  double nx = row[7 + 3 * b], ny = row[7 + 3 * b + 1], nz = row[7 + 3 * b + 2];
  for (int i = 0; i < T.K; ++i) {
    const double x = nx, y = ny, z = nz;
    const int bi = b;
    if (i + 1 < T.K) {
      b = T.order[i + 1];
      nx = row[7 + 3 * b]; ny = row[7 + 3 * b + 1]; nz = row[7 + 3 * b + 2];
    }
    const int s = T.src[i];
    if (s >= 0) {
      HRG_UNROLL
      for (int k = 0; k < kHumanSlots; ++k)
        if (s == k) {
          HRG_UNROLL
          for (int m = 0; m < 12; ++m) cur[m] = slot[k][m];
        }
    }
    const double rx = T.rel[i][0], ry = T.rel[i][1], rz = T.rel[i][2];
    const double ax = cur[9] + fma(cur[0], rx, fma(cur[1], ry, cur[2] * rz));
    const double ay = cur[10] + fma(cur[3], rx, fma(cur[4], ry, cur[5] * rz));
    const double az = cur[11] + fma(cur[6], rx, fma(cur[7], ry, cur[8] * rz));
    kp[3 * bi] = ax; kp[3 * bi + 1] = ay; kp[3 * bi + 2] = az;
    if (T.has_child[i]) {
      double Rl[9], Rw[9];
      hinge_zyx(x, y, z, Rl);
      mat3_mul(cur, Rl, Rw);
      HRG_UNROLL
      for (int m = 0; m < 9; ++m) cur[m] = Rw[m];
      cur[9] = ax; cur[10] = ay; cur[11] = az;
      const int sv = T.save[i];
      if (sv >= 0) {
        HRG_UNROLL
        for (int k = 1; k < kHumanSlots; ++k)
          if (sv == k) {
            HRG_UNROLL
            for (int m = 0; m < 12; ++m) slot[k][m] = cur[m];
          }
      }
    }
  }
}

// One raw frame row -> one baked row [3 + 3K]: u = Rc (p + offset), then every site walked from a root at the origin
// with rotation Rc Ra. clip_tab [12] = position_offset, Rc.
HRG_HD void human_bake_row(const HumanTreeDev& T, const double* __restrict__ row, const double* __restrict__ clip_tab,
                           double* __restrict__ baked) {
  const double px = row[0] + clip_tab[0], py = row[1] + clip_tab[1], pz = row[2] + clip_tab[2];
  const double* Rc = clip_tab + 3;
  baked[0] = Rc[0] * px + Rc[1] * py + Rc[2] * pz;
  baked[1] = Rc[3] * px + Rc[4] * py + Rc[5] * pz;
  baked[2] = Rc[6] * px + Rc[7] * py + Rc[8] * pz;
  double Ra[9], root[12];
  quat_xyzw_to_mat(row[3], row[4], row[5], row[6], Ra);  // Pelvis_quat is (x,y,z,w)
  mat3_mul(Rc, Ra, root);
  root[9] = root[10] = root[11] = 0.0;
  human_tree_walk(T, row, root, baked + 3);
}

// placement [7] = human_pos_offset, human_rot_offset (w,x,y,z) -> [12] = human_pos_offset, Rp
HRG_HD void human_placement(const double* __restrict__ p7, double* __restrict__ out) {
  out[0] = p7[0]; out[1] = p7[1]; out[2] = p7[2];
  quat_xyzw_to_mat(p7[4], p7[5], p7[6], p7[3], out + 3);
}

// coordinate `axis` of one keypoint: P [12] = human_pos_offset + u, Rp; w = the body's baked site
HRG_HD double human_apply(const double* P, const double* w, int axis) {
  const double* R = P + 3 + 3 * axis;
  return P[axis] + fma(R[0], w[0], fma(R[1], w[1], R[2] * w[2]));
}

#if defined(__CUDACC__)
// `_control_human` time management of one env (human_env.py:1718-1731): returns the bank row, updates the replay state.
__device__ inline int64_t human_pick_frame(const HumanDev& H, int e, int64_t low_level_time, int* clip_o, int* frame_o) {
  const int64_t control_time = (int64_t)floor((double)low_level_time / H.step_length);
  int idx = H.list_idx[e];
  int64_t start = H.anim_start[e];
  int64_t at = control_time - start;
  int clip = H.clip_list[(size_t)e * H.M + idx];
  if (at > (int64_t)H.clip_len[clip] - 1) {  // _progress_to_next_animation (human_env.py:1700-1708)
    idx = (idx + 1) % H.M;
    at = 0;
    start = control_time;
    clip = H.clip_list[(size_t)e * H.M + idx];
    H.list_idx[e] = idx;
    H.anim_start[e] = start;
  }
  if (at < 0) at = 0;  // a clock behind the env's start time has no frame: show the first one
  *clip_o = clip;
  *frame_o = (int)at;
  return H.clip_off[clip] + at;
}

// Once per clip bank: lane = frame row.
__global__ void __launch_bounds__(128) k_human_bake(HumanTreeDev T, int C, int64_t n_rows, const double* __restrict__ frames,
                                                    const int32_t* __restrict__ row_clip,
                                                    const double* __restrict__ clip_tab, double* __restrict__ baked) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rows) return;
  human_bake_row(T, frames + r * C, clip_tab + (size_t)row_clip[r] * 12, baked + r * (3 + 3 * T.K));
}

// Every tick: block = kHumanTileEnvs consecutive envs.
__global__ void __launch_bounds__(kHumanBlock) k_human_keypoints(HumanDev H, const int64_t* __restrict__ low_level_time,
                                                                double* __restrict__ keypoints,
                                                                int32_t* __restrict__ clip_out,
                                                                int32_t* __restrict__ frame_out) {
  __shared__ double s_P[kHumanTileEnvs][13];  // human_pos_offset + u, Rp (odd stride)
  __shared__ const double* s_w[kHumanTileEnvs];
  const int e0 = blockIdx.x * kHumanTileEnvs;
  const int n_live = min(kHumanTileEnvs, H.N - e0);
  const int KO = 3 * H.K;
  if (threadIdx.x < n_live) {
    const int e = e0 + threadIdx.x;
    int clip = 0, frame = 0;
    const int64_t rr = human_pick_frame(H, e, low_level_time[e], &clip, &frame);
    if (clip_out) clip_out[e] = clip;
    if (frame_out) frame_out[e] = frame;
    const double* __restrict__ row = H.baked + rr * (3 + KO);
    const double* __restrict__ pl = H.placement + (size_t)e * 12;
    double* P = s_P[threadIdx.x];
    HRG_UNROLL
    for (int i = 0; i < 3; ++i) P[i] = pl[i] + row[i];
    HRG_UNROLL
    for (int i = 3; i < 12; ++i) P[i] = pl[i];
    s_w[threadIdx.x] = row + 3;
  }
  __syncthreads();
  double* __restrict__ dst = keypoints + (size_t)e0 * KO;
  for (int i = threadIdx.x; i < n_live * KO; i += kHumanBlock) {
    const int el = i / KO, c = i - el * KO, b = c / 3, axis = c - 3 * b;
    dst[i] = human_apply(s_P[el], s_w[el] + 3 * b, axis);
  }
}

__global__ void k_human_reset(HumanDev H, const uint8_t* __restrict__ mask, const double* __restrict__ placement,
                              const int32_t* __restrict__ clip_list) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= H.N) return;
  if (mask && !mask[e]) return;
  human_placement(placement + (size_t)e * 7, H.placement + (size_t)e * 12);
  for (int i = 0; i < H.M; ++i) {  // ids outside the bank are clamped (the host layer rejects them before they get here)
    const int c = clip_list[(size_t)e * H.M + i];
    H.clip_list[(size_t)e * H.M + i] = c < 0 ? 0 : (c >= H.n_clips ? H.n_clips - 1 : c);
  }
  H.list_idx[e] = 0;
  H.anim_start[e] = 0;
}
#endif  // __CUDACC__

}  // namespace hrg
