// Capsule geometry of the shield: ReachLib capsule-capsule distance and the human Pos/Vel/Accel occupancy models.
// RL = SS/external/SaRA/ReachLib (reference tree). Branch structure follows the reference literally;
// the translation unit is compiled with -fmad=false so that every add/mul/div/sqrt rounds like the x86-64 build.
#pragma once
#include "path_math.cuh"

namespace hrg {

struct V3 {
  double x, y, z;
};
HRG_HD V3 v3(double x, double y, double z) {
  V3 r;
  r.x = x; r.y = y; r.z = z;
  return r;
}
HRG_HD V3 vsub(const V3& a, const V3& b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
HRG_HD V3 vadd(const V3& a, const V3& b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
HRG_HD V3 vscale(const V3& a, double s) { return v3(a.x * s, a.y * s, a.z * s); }
HRG_HD bool veq(const V3& a, const V3& b) { return a.x == b.x && a.y == b.y && a.z == b.z; }  // RL point.hpp:93-99
HRG_HD double vnorm(const V3& p) { return sqrt(p.x * p.x + p.y * p.y + p.z * p.z); }          // RL point.cpp:34-36
// vnorm for operands that are often exactly zero (a resting link, two balls with one centre): the device sqrt sends 0
// down its slow special-case path, lane by lane; sqrt(1) + a select of the exact result 0 stays on the fast path.
HRG_HD double vnorm_z(const V3& p) {
  const double s = p.x * p.x + p.y * p.y + p.z * p.z;
  const double r = sqrt(s == 0.0 ? 1.0 : s);
  return s == 0.0 ? 0.0 : r;
}
HRG_HD double vdot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }    // RL point.cpp:42-44
HRG_HD V3 vcross(const V3& a, const V3& b) {                                                   // RL point.cpp:46-52
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
HRG_HD double vdet3(const V3& X, const V3& Y, const V3& Z) {  // RL point.cpp:54-60
  return X.x * (Y.y * Z.z - Y.z * Z.y) - X.y * (Y.x * Z.z - Y.z * Z.x) + X.z * (Y.x * Z.y - Y.y * Z.x);
}

struct Cap {
  V3 p1, p2;
  double r;
};
HRG_HD Cap cap_load(const double* c) {
  Cap k;
  k.p1 = v3(c[0], c[1], c[2]);
  k.p2 = v3(c[3], c[4], c[5]);
  k.r = c[6];
  return k;
}
HRG_HD void cap_store(const Cap& k, double* c) {
  c[0] = k.p1.x; c[1] = k.p1.y; c[2] = k.p1.z;
  c[3] = k.p2.x; c[4] = k.p2.y; c[5] = k.p2.z;
  c[6] = k.r;
}

// Axis of a capsule: d = p2 - p1, its length and the unit vector d / |d| (NaN for a sphere, where it is never used).
// Every distance routine below needs these of both capsules; k_reach_verify evaluates J x H pairs per env, so it
// derives them once per link and once per human capsule instead of once per pair - same operations, same bits.
struct Axis {
  V3 d;
  double mag;
  V3 u;
};
HRG_HD Axis axis_of(const Cap& c) {
  Axis a;
  a.d = vsub(c.p2, c.p1);
  a.mag = 0.0;
  a.u = v3(0, 0, 0);
  if (veq(c.p1, c.p2)) return a;  // sphere: no routine reads its axis, and 0 / 0 would only visit the slow division path
  a.mag = vnorm(a.d);
  a.u = v3(a.d.x / a.mag, a.d.y / a.mag, a.d.z / a.mag);
  return a;
}

// point_line_segment_dist (RL/include/intersections.hpp:62-84)
HRG_HD double point_segment_dist(const V3& p, const Cap& c, const Axis& ax) {
  const V3& se = ax.d;
  double dse = ax.mag;
  double dsp = vnorm(vsub(p, c.p1));
  V3 ep = vsub(p, c.p2);
  double dep = vnorm(ep);
  if (sqrt(dse * dse + dep * dep) <= dsp) return dep;
  if (sqrt(dse * dse + dsp * dsp) <= dep) return dsp;
  return vnorm(vcross(se, ep)) / dse;
}
HRG_HD double point_segment_dist(const V3& p, const Cap& c) { return point_segment_dist(p, c, axis_of(c)); }

// The parallel-segments branch of min_segment_distance (RL/src/intersections.cpp:80-104), including its chained
// comparisons: `d0 <= 0 >= d1` parses as ((d0 <= 0) ? 1.0 : 0.0) >= d1, `d0 >= magA <= d1` as
// ((d0 >= magA) ? 1.0 : 0.0) <= d1. Rare (exactly parallel axes); inlined all the same - as a real call its by-reference operands were spilled to the stack for every pair.
HRG_HD double parallel_segment_distance(const Cap& c1, const Axis& a1, const Cap& c2) {
  const V3& _A = a1.u;
  const double magA = a1.mag;
  double d0 = vdot(_A, vsub(c2.p1, c1.p1));
  double d1 = vdot(_A, vsub(c2.p2, c1.p1));
  if (((d0 <= 0) ? 1.0 : 0.0) >= d1) {
    if (fabs(d0) < fabs(d1)) return vnorm(vsub(c1.p1, c2.p1));
    return vnorm(vsub(c1.p1, c2.p2));
  } else if (((d0 >= magA) ? 1.0 : 0.0) <= d1) {
    if (fabs(d0) < fabs(d1)) return vnorm(vsub(c1.p2, c2.p1));
    return vnorm(vsub(c1.p2, c2.p2));
  }
  return vnorm(vsub(vadd(v3(_A.x * d0, _A.y * d0, _A.z * d0), c1.p1), c2.p1));
}

HRG_HD V3 vsel(bool k, const V3& a, const V3& b) { return v3(k ? a.x : b.x, k ? a.y : b.y, k ? a.z : b.z); }

// min_segment_distance (RL/src/intersections.cpp:65-140). The clamping cascade of the reference (:112-137) is written
// with selects: 32 envs share a warp and nearly every warp sees every case, so straight-line code that computes both
// re-projections and keeps the one the reference would have taken is cheaper than the divergent branches. The values
// selected are the ones the reference computes, operation for operation.
HRG_HD double min_segment_distance(const Cap& c1, const Axis& a1, const Cap& c2, const Axis& a2) {
  const V3& _A = a1.u;
  const V3& _B = a2.u;
  const double magA = a1.mag, magB = a2.mag;
  V3 cross = vcross(_A, _B);
  double nc = vnorm(cross);
  double denom = nc * nc;  // std::pow(norm, 2)
  if (denom == 0) return parallel_segment_distance(c1, a1, c2);
  V3 t = vsub(c2.p1, c1.p1);
  double detA = vdet3(t, _B, cross);
  double detB = vdet3(t, _A, cross);
  double t0 = detA / denom, t1 = detB / denom;
  V3 pA = vadd(c1.p1, v3(_A.x * t0, _A.y * t0, _A.z * t0));
  V3 pB = vadd(c2.p1, v3(_B.x * t1, _B.y * t1, _B.z * t1));
#ifdef HRG_BRANCHY_CLAMP
  if (t0 < 0)
    pA = c1.p1;
  else if (t0 > magA)
    pA = c1.p2;
  else if (t1 < 0)
    pB = c2.p1;
  else if (t1 > magB)
    pB = c2.p2;
  if (t0 < 0 || t0 > magA) {
    double dot = vdot(_B, vsub(pA, c2.p1));
    if (dot < 0)
      dot = 0;
    else if (dot > magB)
      dot = magB;
    pB = vadd(c2.p1, v3(_B.x * dot, _B.y * dot, _B.z * dot));
  }
  if (t1 < 0 || t1 > magB) {
    double dot = vdot(_A, vsub(pB, c1.p1));
    if (dot < 0)
      dot = 0;
    else if (dot > magA)
      dot = magA;
    pA = vadd(c1.p1, v3(_A.x * dot, _A.y * dot, _A.z * dot));
  }
#else
  const bool a_lo = t0 < 0, a_hi = t0 > magA, a_out = a_lo || a_hi;
  const bool b_lo = t1 < 0, b_hi = t1 > magB, b_out = b_lo || b_hi;
  // if (t0 < 0) pA = p1A; else if (t0 > magA) pA = p2A; else if (t1 < 0) pB = p1B; else if (t1 > magB) pB = p2B;
  pA = vsel(a_lo, c1.p1, vsel(a_hi, c1.p2, pA));
  pB = vsel(!a_out && b_lo, c2.p1, vsel(!a_out && b_hi, c2.p2, pB));
  {  // if (t0 < 0 || t0 > magA): project pA onto segment B
    double dot = vdot(_B, vsub(pA, c2.p1));
    dot = dot < 0 ? 0.0 : (dot > magB ? magB : dot);
    pB = vsel(a_out, vadd(c2.p1, v3(_B.x * dot, _B.y * dot, _B.z * dot)), pB);
  }
  {  // if (t1 < 0 || t1 > magB): project pB onto segment A
    double dot = vdot(_A, vsub(pB, c1.p1));
    dot = dot < 0 ? 0.0 : (dot > magA ? magA : dot);
    pA = vsel(b_out, vadd(c1.p1, v3(_A.x * dot, _A.y * dot, _A.z * dot)), pA);
  }
#endif
  return vnorm(vsub(pA, pB));
}
HRG_HD double min_segment_distance(const Cap& c1, const Cap& c2) {
  return min_segment_distance(c1, axis_of(c1), c2, axis_of(c2));
}

// capsule_capsule_dist (RL/include/intersections.hpp:105-124); collision iff the result is < 0.0 (:129-131).
HRG_HD double capsule_capsule_dist(const Cap& c1, const Axis& a1, const Cap& c2, const Axis& a2) {
  bool s1 = veq(c1.p1, c1.p2), s2 = veq(c2.p1, c2.p2);
  if (s1 && s2) {
    V3 d = vsub(c1.p1, c2.p1);
    return vnorm(d) - c1.r - c2.r;
  } else if (s1 && !s2) {
    return point_segment_dist(c1.p1, c2, a2) - c1.r - c2.r;
  } else if (!s1 && s2) {
    return point_segment_dist(c2.p1, c1, a1) - c1.r - c2.r;
  }
  return min_segment_distance(c1, a1, c2, a2) - c1.r - c2.r;
}
HRG_HD double capsule_capsule_dist(const Cap& c1, const Cap& c2) {
  return capsule_capsule_dist(c1, axis_of(c1), c2, axis_of(c2));
}

// ---- human occupancy models --------------------------------------------------------------------------------------
struct Ball {
  V3 p;
  double r;
};

// Capsule::ballEnclosure (RL/src/capsule.cpp:82-109) with alphaBeta (:122-125) and pk (:111-120)
HRG_HD Ball ball_enclosure(const Ball& c1, const Ball& c2) {
  double ri, rj;
  V3 pi, pj;
  if (c1.r >= c2.r) {
    ri = c1.r; rj = c2.r; pi = c1.p; pj = c2.p;
  } else {
    ri = c2.r; rj = c1.r; pi = c2.p; pj = c1.p;
  }
  V3 x = vsub(pi, pj);
  double n = vnorm_z(x);  // the Vel model encloses two balls with the SAME centre
  double alpha = fmax(ri - rj, n), beta = fmin(ri - rj, n);
  V3 pk = pj;
  if (!(n == 0.0)) pk = v3(pj.x + (x.x / n) * beta, pj.y + (x.y / n) * beta, pj.z + (x.z / n) * beta);
  Ball out;
  out.p = v3((pi.x + pk.x) / 2.0, (pi.y + pk.y) / 2.0, (pi.z + pk.z) / 2.0);
  out.r = (ri + rj + alpha) / 2.0;
  return out;
}

// Extremity::update (RL/src/extremity.cpp:39-56) via ArticulatedPos::update (t_a = 0).
HRG_HD Cap human_pos_capsule(const V3& base, double v_max, double length, double thickness, double t_b, double delay,
                             double e_pos) {
  double delta_t = t_b - 0.0 + delay;
  double dist = v_max * delta_t;
  Cap c;
  c.p1 = base;
  c.p2 = base;
  c.r = dist + length + thickness / 2.0 + e_pos;
  return c;
}

// Shared tail of BodyPartVel::update / BodyPartAccel::update (RL/src/body_part_vel.cpp:42-90,
// body_part_accel.cpp:37-95): b1/b2 are the ball enclosures of the proximal/distal joint over [t_a, t_b]; thickness/2
// goes to the larger one; capsuleEnclosure (RL/src/capsule.cpp:56-80) then returns (b1.p, b2.p, max r).
HRG_HD Cap body_capsule(const Ball& b1_in, const Ball& b2_in, double thickness) {
  Ball b1 = b1_in, b2 = b2_in;
  if (b1.r >= b2.r)
    b1.r += thickness / 2.0;
  else
    b2.r += thickness / 2.0;
  Cap c;
  c.p1 = b1.p;
  c.p2 = b2.p;
  c.r = (b1.r >= b2.r) ? b1.r : b2.r;
  return c;
}

// BodyPartVel (RL/src/body_part_vel.cpp:42-103): ry = Capsule(p, p, e_pos + max_v * (t + delay)), t_a = 0.
HRG_HD Cap human_vel_capsule(const V3& p1, const V3& p2, double vmax1, double vmax2, double thickness, double t_b,
                             double delay, double e_pos) {
  Ball a1, a2;
  a1.p = p1; a1.r = e_pos + vmax1 * (0.0 + delay);
  a2.p = p1; a2.r = e_pos + vmax1 * (t_b + delay);
  if (veq(p1, p2)) {  // ball: occupancy = rp1_t2 with r += thickness/2
    Cap c;
    c.p1 = a2.p; c.p2 = a2.p; c.r = a2.r + thickness / 2.0;
    return c;
  }
  Ball b1 = ball_enclosure(a1, a2);
  Ball d1, d2;
  d1.p = p2; d1.r = e_pos + vmax2 * (0.0 + delay);
  d2.p = p2; d2.r = e_pos + vmax2 * (t_b + delay);
  Ball b2 = ball_enclosure(d1, d2);
  return body_capsule(b1, b2, thickness);
}

// BodyPartAccel::ry (RL/src/body_part_accel.cpp:97-119): centre y + dy * t_total,
// r = (e_pos + e_vel * t_total) + a_max / 2 * t_total^2 (two Minkowski sums, in this order).
HRG_HD Ball accel_ball(const V3& p, const V3& v, double a_max, double t, double delay, double e_pos, double e_vel) {
  double t_total = t + delay;
  Ball b;
  V3 dy = v3(v.x * t_total, v.y * t_total, v.z * t_total);
  V3 c = vadd(p, dy);
  b.p = v3(c.x + 0.0, c.y + 0.0, c.z + 0.0);
  b.r = (e_pos + e_vel * t_total) + a_max / 2.0 * (t_total * t_total);
  return b;
}

// BodyPartAccel::update (RL/src/body_part_accel.cpp:37-95), t_a = 0.
HRG_HD Cap human_accel_capsule(const V3& p1, const V3& p2, const V3& v1, const V3& v2, double amax1, double amax2,
                               double thickness, double t_b, double delay, double e_pos, double e_vel) {
  Ball a1 = accel_ball(p1, v1, amax1, 0.0, delay, e_pos, e_vel);
  Ball a2 = accel_ball(p1, v1, amax1, t_b, delay, e_pos, e_vel);
  if (veq(p1, p2)) {
    Cap c;
    c.p1 = a2.p; c.p2 = a2.p; c.r = a2.r + thickness / 2.0;
    return c;
  }
  Ball b1 = ball_enclosure(a1, a2);
  Ball d1 = accel_ball(p2, v2, amax2, 0.0, delay, e_pos, e_vel);
  Ball d2 = accel_ball(p2, v2, amax2, t_b, delay, e_pos, e_vel);
  Ball b2 = ball_enclosure(d1, d2);
  return body_capsule(b1, b2, thickness);
}

// Human capsule h of the set {Pos[E], Vel[B], Accel[B]} (HumanReach::getAllCapsules, SS/include/safety_shield/
// human_reach.h:203-205) from the measured keypoints; t_reach as in human_reach.cc:76-80. Coordinate `comp` of keypoint
// `k` is kp[(3 * k + comp) * ks]: ks = 1 for one env's contiguous [K*3] record, ks = N (with kp pointing at the env's
// column) for the field-major device state.
HRG_HD V3 kp_point(const double* kp, int k, size_t ks) {
  return v3(kp[(size_t)(3 * k) * ks], kp[(size_t)(3 * k + 1) * ks], kp[(size_t)(3 * k + 2) * ks]);
}
HRG_HD Cap human_capsule(const HrgConfig& c, int h, const double* kp, const double* kv, size_t ks, double t_reach) {
  const int E = c.n_ext, B = c.n_bodies;
  if (h < E)
    return human_pos_capsule(kp_point(kp, c.ext_base[h], ks), c.ext_vmax[h], c.ext_length[h], c.ext_thickness[h],
                             t_reach, c.delay, c.meas_err_pos);
  const int b = (h - E) % B;
  const bool accel = (h - E) >= B;
  const V3 p1 = kp_point(kp, c.body_prox[b], ks), p2 = kp_point(kp, c.body_dist[b], ks);
  if (!accel)
    return human_vel_capsule(p1, p2, c.body_vmax[b][0], c.body_vmax[b][1], c.body_thickness[b], t_reach, c.delay,
                             c.meas_err_pos);
  return human_accel_capsule(p1, p2, kp_point(kv, c.body_prox[b], ks), kp_point(kv, c.body_dist[b], ks),
                             c.body_amax[b][0], c.body_amax[b][1], c.body_thickness[b], t_reach, c.delay,
                             c.meas_err_pos, c.meas_err_vel);
}

// Rows 0..2 of the base pose matrix from (x, y, z, roll, pitch, yaw) — RobotReach ctor / reset
// (SS/src/robot_reach.cc:5-17,41-52). out[12] row-major 3x4.
HRG_HD void base_pose_matrix(double x, double y, double z, double roll, double pitch, double yaw, double* out) {
  double cr = cos(roll), sr = sin(roll), cp = cos(pitch), sp = sin(pitch), cy = cos(yaw), sy = sin(yaw);
  out[0] = cr * cp; out[1] = cr * sp * sy - sr * cy; out[2] = cr * sp * cy + sr * sy; out[3] = x;
  out[4] = sr * cp; out[5] = sr * sp * sy + cr * cy; out[6] = sr * sp * cy - cr * sy; out[7] = y;
  out[8] = -sp;     out[9] = cp * sy;                out[10] = cp * cy;               out[11] = z;
}

// model index (0 Pos, 1 Vel, 2 Accel) of human capsule h
HRG_HD int human_model_of(const HrgConfig& c, int h) { return h < c.n_ext ? 0 : ((h - c.n_ext) < c.n_bodies ? 1 : 2); }

// PlotCapsule.update_pos (HRG/controllers/failsafe_controller/failsafe_controller/plot_capsule.py:30-86): a capsule
// (p1, p2, r) as the pos / size / mat triple of a MuJoCo marker - centre, (r, r, half length), and the rotation that takes
// the z axis onto the capsule axis (shortest-arc quaternion, normalised twice as the reference and scipy do).
HRG_HD void capsule_marker(const double* c, double* pos, double* size, double* mat) {
  const double vx = c[3] - c[0], vy = c[4] - c[1], vz = c[5] - c[2];
  const double norm = sqrt(vx * vx + vy * vy + vz * vz);
  pos[0] = (c[0] + c[3]) / 2; pos[1] = (c[1] + c[4]) / 2; pos[2] = (c[2] + c[5]) / 2;
  size[0] = c[6]; size[1] = c[6];
  double x = 0.0, y = 0.0, z = 0.0, w = 1.0;
  if (norm > 1e-6) {
    size[2] = norm / 2;
    const double ax = -vy / norm, ay = vx / norm, aw = 1 + vz / norm;
    const double nq = sqrt(aw * aw + ax * ax + ay * ay);
    if (nq > 1e-6) {
      x = ax / nq; y = ay / nq; w = aw / nq;
      const double n2 = sqrt(x * x + y * y + z * z + w * w);  // Rotation.from_quat normalises again
      x /= n2; y /= n2; w /= n2;
    }
  } else {
    size[2] = 0.00001;
  }
  // scipy Rotation.as_matrix
  const double x2 = x * x, y2 = y * y, z2 = z * z, w2 = w * w, xy = x * y, zw = z * w, xz = x * z, yw = y * w, yz = y * z,
               xw = x * w;
  mat[0] = x2 - y2 - z2 + w2; mat[1] = 2 * (xy - zw);       mat[2] = 2 * (xz + yw);
  mat[3] = 2 * (xy + zw);     mat[4] = -x2 + y2 - z2 + w2;  mat[5] = 2 * (yz - xw);
  mat[6] = 2 * (xz - yw);     mat[7] = 2 * (yz + xw);       mat[8] = -x2 - y2 + z2 + w2;
}

}  // namespace hrg
