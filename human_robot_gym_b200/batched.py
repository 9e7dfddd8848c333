"""BatchedSafetyShield — N independent sara-shield instances stepped together on one B200.

Same verbs as the reference `safety_shield_py.SafetyShield` (SS/bindings/safety_shield_bindings.cpp:82-115) over
`[N, ...]` FP64 CUDA tensors. PyTorch is only plumbing here (device memory, streams); the arithmetic runs in the
hand-written sm_100a kernels of libhrg_shield.so behind the C ABI of include/hrg_shield.h. No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from . import cabi
from .config import ShieldConfig


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


class BatchedSafetyShield:
    def __init__(self, n_envs: int, sample_time: float, trajectory_config_file: str, robot_config_file: str,
                 mocap_config_file: str, base_pose=None, init_qpos=None, shield_type=cabi.SHIELD_SSM,
                 device: Optional[torch.device] = None, ltt_capacity: Optional[int] = None, store_capsules: bool = False,
                 config: Optional[ShieldConfig] = None):
        """ltt_capacity: samples per long-term-trajectory slot (two slots per env, 32 J bytes per sample). None = the bound
        of cabi.ltt_capacity_bound(): no plan inside the joint limits can be longer, so behaviour equals the reference's
        unbounded LTT. A smaller explicit capacity saves HBM (a FailsafeController goal is at most 0.2 rad away, ~230
        samples); a plan that does not fit is then rejected (HRG_ST_LTT_TRUNCATED) and counted in
        stats()["n_ltt_truncated"] - see check_health()."""
        self.lib = cabi.load_library()  # raises if the CUDA extension is missing
        if not torch.cuda.is_available():
            raise cabi.ShieldLibraryError("BatchedSafetyShield needs a CUDA device; there is no CPU fallback")
        self.config = config or ShieldConfig.from_files(sample_time, trajectory_config_file, robot_config_file,
                                                        mocap_config_file)
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.n_envs = int(n_envs)
        if ltt_capacity is None:
            ltt_capacity = cabi.ltt_capacity_bound(self.config)
        self.ltt_capacity = int(ltt_capacity)
        self.cfg = cabi.make_cfg(self.config, self.ltt_capacity)
        self.J, self.K = self.cfg.nb_joints, self.cfg.n_keypoints
        self.H = self.cfg.n_ext + 2 * self.cfg.n_bodies
        N, J = self.n_envs, self.J
        pose = np.zeros((N, 6)) if base_pose is None else np.broadcast_to(np.asarray(base_pose, np.float64), (N, 6))
        q0 = np.zeros((N, J)) if init_qpos is None else np.broadcast_to(np.asarray(init_qpos, np.float64), (N, J))
        ty = np.broadcast_to(np.asarray(shield_type, np.int32), (N,))
        pose, q0, ty = np.ascontiguousarray(pose), np.ascontiguousarray(q0), np.ascontiguousarray(ty)
        h = C.c_void_p()
        dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        cabi.check(self.lib.hrg_shield_create(C.byref(self.cfg), N, dev_index, pose.ctypes.data, q0.ctypes.data,
                                              ty.ctypes.data, C.byref(h)), "hrg_shield_create")
        self._h = h
        f64 = dict(dtype=torch.float64, device=self.device)
        self.q = torch.zeros(N, J, **f64)
        self.dq = torch.zeros(N, J, **f64)
        self.ddq = torch.zeros(N, J, **f64)
        self.dddq = torch.zeros(N, J, **f64)
        self.is_safe = torch.zeros(N, dtype=torch.uint8, device=self.device)
        self.branch = torch.zeros(N, dtype=torch.uint8, device=self.device)
        self.status = torch.zeros(N, dtype=torch.int32, device=self.device)
        self._stats = torch.zeros(cabi.STATS_WORDS, dtype=torch.int64, device=self.device)
        self._store = False
        if store_capsules:
            self.store_capsules(True)

    # ------------------------------------------------------------------------------------------------ helpers
    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _f64(self, x, shape):
        t = torch.as_tensor(x, dtype=torch.float64, device=self.device)
        return t.expand(shape).contiguous() if t.shape != torch.Size(shape) else t.contiguous()

    def _mask(self, m):
        if m is None:
            return None
        return torch.as_tensor(m, device=self.device).to(torch.uint8).contiguous()

    def close(self):
        if getattr(self, "_h", None):
            torch.cuda.synchronize(self.device)
            self.lib.hrg_shield_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---------------------------------------------------------------------------------------------------- verbs
    def reset(self, base_pose, init_qpos, current_time=0.0, shield_type=cabi.SHIELD_SSM, env_mask=None):
        N, J = self.n_envs, self.J
        pose, q0 = self._f64(base_pose, (N, 6)), self._f64(init_qpos, (N, J))
        t = self._f64(current_time, (N,))
        ty = None  # NULL = all SSM
        if isinstance(shield_type, torch.Tensor):
            ty = shield_type.to(device=self.device, dtype=torch.int32).expand(N).contiguous()
            any_pfl = bool((ty == cabi.SHIELD_PFL).any().item())
        else:
            st = np.asarray(shield_type)
            any_pfl = bool(np.any(st == cabi.SHIELD_PFL))
            if np.any(st != cabi.SHIELD_SSM):
                ty = torch.as_tensor(st.astype(np.int32), device=self.device).expand(N).contiguous()
        if any_pfl:  # PFL needs the per-sample Cartesian velocity buffers; allocate them before the first PFL reset
            cabi.check(self.lib.hrg_shield_enable_pfl(self._h), "hrg_shield_enable_pfl")
        m = self._mask(env_mask)
        cabi.check(self.lib.hrg_shield_reset(self._h, _ptr(m), _ptr(pose), _ptr(q0), _ptr(t), _ptr(ty), self._stream()),
                   "hrg_shield_reset")

    def humanMeasurement(self, keypoints, time, env_mask=None):
        kp = self._f64(keypoints, (self.n_envs, self.K, 3))
        t = self._f64(time, (self.n_envs,))
        m = self._mask(env_mask)
        cabi.check(self.lib.hrg_shield_human_measurement(self._h, _ptr(m), _ptr(kp), _ptr(t), self._stream()),
                   "hrg_shield_human_measurement")

    def newLongTermTrajectory(self, goal_position, goal_velocity=None, env_mask=None):
        gq = self._f64(goal_position, (self.n_envs, self.J))
        gd = None if goal_velocity is None else self._f64(goal_velocity, (self.n_envs, self.J))
        m = self._mask(env_mask)
        cabi.check(self.lib.hrg_shield_new_goal(self._h, _ptr(m), _ptr(gq), _ptr(gd), self._stream()),
                   "hrg_shield_new_goal")

    def step(self, cycle_begin_time):
        """One shield tick for every env. Returns (q, dq, ddq, dddq) views of the output tensors [N, J]."""
        t = self._f64(cycle_begin_time, (self.n_envs,))
        cabi.check(self.lib.hrg_shield_step(self._h, _ptr(t), _ptr(self.q), _ptr(self.dq), _ptr(self.ddq),
                                            _ptr(self.dddq), _ptr(self.is_safe), _ptr(self.branch), _ptr(self.status),
                                            self._stream()), "hrg_shield_step")
        return self.q, self.dq, self.ddq, self.dddq

    def getSafety(self):
        return self.is_safe.bool()

    def store_capsules(self, on=True):
        cabi.check(self.lib.hrg_shield_store_capsules(self._h, int(bool(on))), "hrg_shield_store_capsules")
        self._store = bool(on)

    def capsules(self):
        """(robot [N,J,7], human [N,H,7], min_dist [N,3]) of the last step; needs store_capsules(True) beforehand."""
        f64 = dict(dtype=torch.float64, device=self.device)
        robot = torch.zeros(self.n_envs, self.J, 7, **f64) if self._store else None
        human = torch.zeros(self.n_envs, self.H, 7, **f64) if self._store else None
        md = torch.zeros(self.n_envs, 3, **f64)
        cabi.check(self.lib.hrg_shield_get_capsules(self._h, _ptr(robot), _ptr(human), _ptr(md), self._stream()),
                   "hrg_shield_get_capsules")
        return robot, human, md

    def set_ltt(self, env: int, traj: np.ndarray, sample_time: float, starting_index: int, alpha_i) -> int:
        traj = np.ascontiguousarray(traj, dtype=np.float64).reshape(-1, 4, self.J)
        alpha = np.ascontiguousarray(np.broadcast_to(np.asarray(alpha_i, np.float64), (self.J,)))
        verdict = C.c_int32(-1)
        torch.cuda.synchronize(self.device)
        cabi.check(self.lib.hrg_shield_set_ltt(self._h, int(env), traj.ctypes.data, traj.shape[0], float(sample_time),
                                               int(starting_index), alpha.ctypes.data, C.addressof(verdict)),
                   "hrg_shield_set_ltt")
        return verdict.value

    def set_ltt_batch(self, env_index, traj, length, sample_time: float, starting_index=None, alpha_i=None) -> torch.Tensor:
        """setLongTermTrajectory for many envs in one launch: traj [M, Lmax, 4, J], length [M], env_index [M] (distinct),
        starting_index [M] or None, alpha_i [M, J] (default alpha_i_max of the config). Returns the verdicts [M] (int32,
        device): 0 installed, 1..4 the reference's exceptions (see set_ltt), -1 env / length out of range."""
        f64 = dict(dtype=torch.float64, device=self.device)
        i32 = dict(dtype=torch.int32, device=self.device)
        traj = torch.as_tensor(traj, **f64).contiguous()
        M, Lmax = traj.shape[0], traj.shape[1]
        if tuple(traj.shape[2:]) != (4, self.J):
            raise ValueError(f"traj must be [M, Lmax, 4, {self.J}], got {tuple(traj.shape)}")
        env_index = torch.as_tensor(env_index, **i32).contiguous()
        length = torch.as_tensor(length, **i32).contiguous()
        start = None if starting_index is None else torch.as_tensor(starting_index, **i32).contiguous()
        if alpha_i is None:
            alpha_i = torch.full((M, self.J), float(self.config.traj.alpha_i_max), **f64)
        alpha = torch.as_tensor(alpha_i, **f64).expand(M, self.J).contiguous()
        verdict = torch.full((M,), -2, **i32)
        with torch.cuda.device(self.device):
            cabi.check(self.lib.hrg_shield_set_ltt_batch(self._h, M, _ptr(env_index), _ptr(traj), Lmax, _ptr(length),
                                                         float(sample_time), _ptr(start), _ptr(alpha), _ptr(verdict),
                                                         self._stream()), "hrg_shield_set_ltt_batch")
        return verdict

    def plot_capsules(self):
        """FailsafeController.get_robot_capsules / get_human_capsules (failsafe_controller.py:383-427) for every env:
        ((pos [N,J,3], size [N,J,3], mat [N,J,3,3]), (pos [N,H,3], size, mat)) of the last step's capsules
        (needs store_capsules=True). Device tensors: the marker buffers a viewer would consume, no viewer here."""
        robot, human, _ = self.capsules()
        out = []
        for caps in (robot, human):
            n = caps.shape[0] * caps.shape[1]
            f64 = dict(dtype=torch.float64, device=self.device)
            pos, size, mat = torch.empty(n, 3, **f64), torch.empty(n, 3, **f64), torch.empty(n, 9, **f64)
            with torch.cuda.device(self.device):
                cabi.check(self.lib.hrg_capsule_markers(n, _ptr(caps.contiguous()), _ptr(pos), _ptr(size), _ptr(mat),
                                                        self._stream()), "hrg_capsule_markers")
            shp = caps.shape[:2]
            out.append((pos.view(*shp, 3), size.view(*shp, 3), mat.view(*shp, 3, 3)))
        return tuple(out)

    def ltt_max_window(self, index=None, window_k: int = 0):
        """getMaxAccelerationWindow(index) / getMaxJerkWindow(index) of every env's live LTT -> ([N,J], [N,J]).
        index: [N] int32 tensor, or None for each env's path_s_discrete_; window_k 0 = the shield's sliding_window_k_."""
        f64 = dict(dtype=torch.float64, device=self.device)
        acc, jerk = torch.empty(self.n_envs, self.J, **f64), torch.empty(self.n_envs, self.J, **f64)
        idx = None if index is None else torch.as_tensor(index, dtype=torch.int32, device=self.device).contiguous()
        with torch.cuda.device(self.device):
            cabi.check(self.lib.hrg_shield_ltt_max_window(self._h, _ptr(idx), int(window_k), _ptr(acc), _ptr(jerk),
                                                          self._stream()), "hrg_shield_ltt_max_window")
        return acc, jerk

    def stats(self, clear=False) -> dict:
        cabi.check(self.lib.hrg_shield_stats(self._h, _ptr(self._stats), int(clear), self._stream()), "hrg_shield_stats")
        v = self._stats.cpu()
        md = v[9:10].view(torch.float64).item()
        return {"n_steps": int(v[0]), "n_unsafe": int(v[1]), "n_branch": [int(v[2]), int(v[3]), int(v[4])],
                "n_replan": int(v[5]), "n_replan_failed": int(v[6]), "n_near_threshold": int(v[7]),
                "ltt_len_sum": int(v[8]), "min_dist": md, "n_ltt_truncated": int(v[10]), "n_pfl_rejected": int(v[11])}

    def check_health(self):
        """Raises if the batch has left the reference's behaviour: a plan longer than ltt_capacity was rejected (the
        reference's LTT is unbounded), or an env asked for PFL on a handle without PFL buffers."""
        st = self.stats()
        if st["n_ltt_truncated"]:
            raise cabi.ShieldLibraryError(
                f"{st['n_ltt_truncated']} replans needed more than ltt_capacity={self.ltt_capacity} samples and were "
                f"rejected; the reference would have accepted them. Use ltt_capacity >= "
                f"{cabi.ltt_capacity_bound(self.config)} (or None)")
        if st["n_pfl_rejected"]:
            raise cabi.ShieldLibraryError(f"{st['n_pfl_rejected']} envs were reset to PFL without hrg_shield_enable_pfl")
        return st

    def stats_tensor(self, clear=False) -> torch.Tensor:
        """Raw HrgStats as an int64 device tensor (word 9 = min_dist bits) for the cross-GPU all_reduce."""
        cabi.check(self.lib.hrg_shield_stats(self._h, _ptr(self._stats), int(clear), self._stream()), "hrg_shield_stats")
        return self._stats

    def state_dict(self) -> dict:
        """Checkpoint of every env's shield state (hrg_shield_state_save): paths, flags, both long-term-trajectory slots,
        the human measurement history, base poses, statistics. Call between ticks; synchronises the device."""
        n = int(self.lib.hrg_shield_state_bytes(self._h))
        blob = torch.empty(n, dtype=torch.uint8)
        with torch.cuda.device(self.device):
            cabi.check(self.lib.hrg_shield_state_save(self._h, blob.data_ptr(), n), "hrg_shield_state_save")
        return {"blob": blob, "n_envs": self.n_envs, "nb_joints": self.J, "ltt_capacity": self.ltt_capacity}

    def load_state_dict(self, state: dict):
        """Restore a state_dict() of a shield of the same shape; the next tick continues bit-identically."""
        blob = state["blob"].contiguous()
        with torch.cuda.device(self.device):
            cabi.check(self.lib.hrg_shield_state_load(self._h, blob.data_ptr(), blob.numel()), "hrg_shield_state_load")

    def profile(self, max_steps: int):
        cabi.check(self.lib.hrg_shield_profile(self._h, int(max_steps)), "hrg_shield_profile")

    def profile_read(self):
        """({kernel: total ms}, steps) of the profiled steps since profile(max_steps)."""
        ms = (C.c_double * 4)()
        n = C.c_int(0)
        cabi.check(self.lib.hrg_shield_profile_read(self._h, C.addressof(ms), C.addressof(n)), "hrg_shield_profile_read")
        return dict(zip(("k_pre_plan", "k_replan", "k_reach_verify", "k_select"), list(ms))), n.value

    def profile_read_kernels(self):
        """({kernel: total ms} for every launch slot of the tick (cabi.PROF_SLOTS), steps) since profile(max_steps)."""
        ms = (C.c_double * len(cabi.PROF_SLOTS))()
        n = C.c_int(0)
        cabi.check(self.lib.hrg_shield_profile_read_kernels(self._h, C.addressof(ms), C.addressof(n)),
                   "hrg_shield_profile_read_kernels")
        return dict(zip(cabi.PROF_SLOTS, list(ms))), n.value

    def launch_count(self) -> int:
        return int(self.lib.hrg_shield_launch_count(self._h))

    def tick_device(self, kp: torch.Tensor, t: torch.Tensor, goal_mask: Optional[torch.Tensor] = None,
                    goal_q: Optional[torch.Tensor] = None):
        """[newLongTermTrajectory on goal_mask] + humanMeasurement + step on already-resident, correctly typed device
        tensors (kp [N,K,3] f64, t [N] f64, goal_mask [N] u8, goal_q [N,J] f64): one ABI call, no conversions."""
        gm = None if goal_mask is None else goal_mask.data_ptr()
        gq = None if goal_mask is None else goal_q.data_ptr()
        cabi.check(self.lib.hrg_shield_tick_device(self._h, kp.data_ptr(), t.data_ptr(), gm, gq, self.q.data_ptr(),
                                                   self.dq.data_ptr(), self.ddq.data_ptr(), self.dddq.data_ptr(),
                                                   self.is_safe.data_ptr(), self.branch.data_ptr(),
                                                   self.status.data_ptr(), self._stream()), "tick_device")

    def tick_host(self, kp_host: torch.Tensor, t_host: torch.Tensor, goal_mask_host, goal_q_host, q_host, dq_host,
                  ddq_host, flags_host):
        """humanMeasurement + [newLongTermTrajectory] + step with pinned HOST tensors (copies on the current stream)."""
        cabi.check(self.lib.hrg_shield_tick_host(self._h, kp_host.data_ptr(), t_host.data_ptr(), _ptr(goal_mask_host),
                                                 _ptr(goal_q_host), _ptr(q_host), _ptr(dq_host), _ptr(ddq_host),
                                                 _ptr(flags_host), self._stream()), "hrg_shield_tick_host")

    def upload_inputs(self, slot: int, kp_host, t_host, goal_mask_host=None, goal_q_host=None, copy_stream=None):
        """Copy one tick's pinned HOST inputs into staging slot 0/1 on `copy_stream` (a torch.cuda.Stream)."""
        cs = (copy_stream or torch.cuda.current_stream(self.device)).cuda_stream
        cabi.check(self.lib.hrg_shield_upload_inputs(self._h, int(slot), kp_host.data_ptr(), t_host.data_ptr(),
                                                     _ptr(goal_mask_host), _ptr(goal_q_host), cs),
                   "hrg_shield_upload_inputs")

    def upload_inputs_sparse(self, slot: int, kp_host, t_host, goal_env_host=None, goal_q_host=None, n_goals=None,
                             copy_stream=None):
        """upload_inputs with the goals as (env index, goal) pairs: goal_env_host [>= n_goals] int32 and goal_q_host
        [>= n_goals, J], pinned HOST tensors; only the envs that start a policy step this tick travel."""
        cs = (copy_stream or torch.cuda.current_stream(self.device)).cuda_stream
        n = 0 if goal_env_host is None else int(goal_env_host.numel() if n_goals is None else n_goals)
        cabi.check(self.lib.hrg_shield_upload_inputs_sparse(self._h, int(slot), kp_host.data_ptr(), t_host.data_ptr(), n,
                                                            _ptr(goal_env_host), _ptr(goal_q_host), cs),
                   "hrg_shield_upload_inputs_sparse")

    def tick_staged(self, slot: int, q_host, dq_host, ddq_host, flags_host):
        """humanMeasurement + [newLongTermTrajectory] + step on a previously uploaded slot; setpoints to pinned HOST."""
        cabi.check(self.lib.hrg_shield_tick_staged(self._h, int(slot), _ptr(q_host), _ptr(dq_host), _ptr(ddq_host),
                                                   _ptr(flags_host), self._stream()), "hrg_shield_tick_staged")


class PipelinedHostShield:
    """Host-buffer front end: the batch is cut into `chunks` independent env blocks, each a BatchedSafetyShield with its
    own CUDA stream, so that inside ONE tick the host->device copy of block i+1, the kernels of block i and the
    device->host copy of block i-1 overlap (separate copy engines + SMs). Envs are independent, so the result is
    identical to one big batch. This is what a vectorised env with host-resident (MuJoCo) state calls once per tick:

        shield.tick_host(kp, t, goal_mask, goal_q, q_out, dq_out, ddq_out, flags_out); shield.synchronize()

    All arguments are pinned host tensors over the WHOLE batch ([N, K, 3], [N], [N] u8, [N, J], 3 x [N, J], [N] u8).
    """

    def __init__(self, n_envs: int, sample_time: float, trajectory_config_file: str, robot_config_file: str,
                 mocap_config_file: str, base_pose=None, init_qpos=None, shield_type=cabi.SHIELD_SSM, device=None,
                 chunks: int = 4, **kw):
        from .sharded import shard_range
        self.n_envs = int(n_envs)
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.bounds = [shard_range(self.n_envs, c, chunks) for c in range(chunks) if shard_range(self.n_envs, c, chunks)[1] >
                       shard_range(self.n_envs, c, chunks)[0]]
        self.streams = [torch.cuda.Stream(device=self.device) for _ in self.bounds]

        def rows(x, lo, hi):
            if x is None:
                return None
            x = np.asarray(x)
            return x[lo:hi] if x.ndim == 2 else x
        ty = np.broadcast_to(np.asarray(shield_type, np.int32), (self.n_envs,))
        self.parts = [BatchedSafetyShield(hi - lo, sample_time, trajectory_config_file, robot_config_file,
                                          mocap_config_file, base_pose=rows(base_pose, lo, hi),
                                          init_qpos=rows(init_qpos, lo, hi), shield_type=ty[lo:hi], device=self.device, **kw)
                      for lo, hi in self.bounds]
        self.J, self.K = self.parts[0].J, self.parts[0].K

    def tick_host(self, kp_host, t_host, goal_mask_host, goal_q_host, q_host, dq_host, ddq_host, flags_host):
        for (lo, hi), st, p in zip(self.bounds, self.streams, self.parts):
            cabi.check(p.lib.hrg_shield_tick_host(
                p._h, kp_host[lo:hi].data_ptr(), t_host[lo:hi].data_ptr(),
                None if goal_mask_host is None else goal_mask_host[lo:hi].data_ptr(),
                None if goal_q_host is None else goal_q_host[lo:hi].data_ptr(),
                _ptr(None if q_host is None else q_host[lo:hi]), _ptr(None if dq_host is None else dq_host[lo:hi]),
                _ptr(None if ddq_host is None else ddq_host[lo:hi]),
                _ptr(None if flags_host is None else flags_host[lo:hi]), st.cuda_stream), "hrg_shield_tick_host")

    def synchronize(self):
        for st in self.streams:
            st.synchronize()

    def launch_count(self) -> int:
        return sum(p.launch_count() for p in self.parts)

    def stats(self, clear=False) -> dict:
        out = None
        for p in self.parts:
            s = p.stats(clear=clear)
            if out is None:
                out = s
            else:
                for k, v in s.items():
                    if k == "min_dist":
                        out[k] = min(out[k], v)
                    elif k == "n_branch":
                        out[k] = [a + b for a, b in zip(out[k], v)]
                    else:
                        out[k] += v
        return out

    def close(self):
        for p in self.parts:
            p.close()
