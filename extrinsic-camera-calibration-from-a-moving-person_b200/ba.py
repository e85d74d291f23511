"""Host side of the bundle-adjustment hot path: mirrors the reference's ba.py API
(to_theta / from_theta / objfun / ba_main, reference ba.py:18-210) on top of libbax.so.

PyTorch is used for device memory, streams and (multi-GPU) torch.distributed only; every
numerical kernel is a hand-written sm_100a CUDA kernel behind include/bax.h.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import numpy as np
import torch

from . import _lib
from ._lib import BaxError, check, lib

# util.OP_BONE (reference util.py:129-145 via OP_KEY util.py:20-46)
OP_BONE = np.array([[8, 1], [2, 5], [2, 3], [5, 6], [4, 3], [7, 6],
                    [9, 8], [12, 8], [9, 10], [10, 11], [12, 13], [13, 14]], dtype=np.int64)
EPS = float(np.finfo(np.float64).eps)


# ----------------------------------------------------------------------------- SO(3) on the host
def rodrigues_batch(rvecs):
    """(n, 3) rvecs -> (n, 3, 3): the map cv2.Rodrigues applies at reference ba.py:29, vectorised (a sweep over
    thousands of problems converts all its cameras at once).  `rodrigues` is this function on one vector, so the
    two agree bit for bit by construction; transcendental functions go through `math` one value at a time (NumPy's
    SIMD cos / sin / arccos need not round like libm's)."""
    rv = np.asarray(rvecs, dtype=np.float64).reshape(-1, 3)
    th = np.sqrt(rv[:, 0] * rv[:, 0] + rv[:, 1] * rv[:, 1] + rv[:, 2] * rv[:, 2])
    small = th < EPS
    k = rv / np.where(small, 1.0, th)[:, None]
    c = np.array([math.cos(x) for x in th]).reshape(-1)
    s = np.array([math.sin(x) for x in th]).reshape(-1)
    kx = np.zeros((len(rv), 3, 3))
    kx[:, 0, 1], kx[:, 0, 2] = -k[:, 2], k[:, 1]
    kx[:, 1, 0], kx[:, 1, 2] = k[:, 2], -k[:, 0]
    kx[:, 2, 0], kx[:, 2, 1] = -k[:, 1], k[:, 0]
    R = c[:, None, None] * np.eye(3)[None] + (1 - c)[:, None, None] * (k[:, :, None] * k[:, None, :]) + s[:, None, None] * kx
    R[small] = np.eye(3)
    return R


def rodrigues(rvec):
    """rvec -> R, the map cv2.Rodrigues applies at reference ba.py:29."""
    return rodrigues_batch(np.asarray(rvec, dtype=np.float64).reshape(1, 3))[0]


def _rodrigues_inv_near_pi(R, th):
    """cv2.Rodrigues' branch for sin(theta) ~ 0 (R already on SO(3))."""
    rx = math.sqrt(max((R[0, 0] + 1) * 0.5, 0.0))
    ry = math.sqrt(max((R[1, 1] + 1) * 0.5, 0.0)) * (-1.0 if R[0, 1] < 0 else 1.0)
    rz = math.sqrt(max((R[2, 2] + 1) * 0.5, 0.0)) * (-1.0 if R[0, 2] < 0 else 1.0)
    if abs(rx) < abs(ry) and abs(rx) < abs(rz) and (R[1, 2] > 0) != (ry * rz > 0):
        rz = -rz
    v = np.array([rx, ry, rz])
    return v * (th / np.linalg.norm(v))


def rodrigues_inv_batch(Rs):
    """(n, 3, 3) -> (n, 3) rvecs as cv2.Rodrigues does at reference ba.py:19: SVD-project on SO(3) first (the linear
    calibration returns non-orthonormal rotations), then the log map.  `rodrigues_inv` is this
    function on one matrix (bit-identical by construction)."""
    Rs = np.asarray(Rs, dtype=np.float64).reshape(-1, 3, 3)
    U, _, Vt = np.linalg.svd(Rs)
    R = U @ Vt
    r = np.stack([R[:, 2, 1] - R[:, 1, 2], R[:, 0, 2] - R[:, 2, 0], R[:, 1, 0] - R[:, 0, 1]], axis=1)
    s = np.sqrt((r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2]) * 0.25)
    c = np.minimum(1.0, np.maximum(-1.0, (R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2] - 1) * 0.5))
    th = np.array([math.acos(x) for x in c]).reshape(-1)
    near = s < 1e-5
    out = r * (0.5 * th / np.where(near, 1.0, s))[:, None]
    for i in np.nonzero(near)[0]:
        out[i] = np.zeros(3) if c[i] > 0 else _rodrigues_inv_near_pi(R[i], th[i])
    return out


def rodrigues_inv(R):
    return rodrigues_inv_batch(np.asarray(R, dtype=np.float64).reshape(1, 3, 3))[0]


def to_theta(R, t, x):
    """reference ba.py:18-20."""
    rv = rodrigues_inv_batch(R).reshape(-1)
    return np.concatenate([rv, np.asarray(t, dtype=np.float64).reshape(-1),
                           np.asarray(x, dtype=np.float64).reshape(-1)])


def from_theta(theta, C):
    """reference ba.py:23-31."""
    theta = np.asarray(theta, dtype=np.float64)
    rv = theta[:3 * C].reshape(C, 3)
    return (rodrigues_batch(rv), theta[3 * C:6 * C].reshape(C, 3, 1), theta[6 * C:].reshape(-1, 3))


# ----------------------------------------------------------------------------- device problem
class _DevView:
    """Zero-copy torch view of a device buffer owned by libbax (for torch.distributed.all_reduce)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f8", "data": (int(ptr), False),
                                         "version": 2}


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


_native_comms = {}


def _native_comm(group, dev):
    """libbax's own NCCL communicator over the ranks of a torch.distributed group (one per group and device, kept for
    the life of the process): rank 0 draws the unique id, torch.distributed hands it round, every rank joins.  The
    trust-region driver in the library then all-reduces the packed buffers itself (bax_allreduce_attach).  Returns None
    when NCCL cannot be bound or BAX_NATIVE_COMM=0 — the host-language all-reduces remain."""
    import torch.distributed as dist
    if os.environ.get("BAX_NATIVE_COMM", "1") == "0" or dist.get_backend(group) != "nccl":
        return None
    # keyed by the group's MEMBERS (global ranks), not by id(group): a destroyed group's id can come back on another object
    key = (tuple(dist.get_process_group_ranks(group)), dev.index)
    if key not in _native_comms:
        L = lib()
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        buf = C.create_string_buffer(128)
        ok = 1
        if rank == 0:
            ok = 1 if L.bax_comm_unique_id(buf) == 0 else 0
        box = [bytes(buf.raw) if ok else None]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0), group=group)
        comm = C.c_void_p()
        good = box[0] is not None and L.bax_comm_create(dev.index or 0, box[0], rank, world, C.byref(comm)) == 0
        flag = torch.tensor([1 if good else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)            # all ranks or none
        if int(flag.item()) == 0:
            if good:
                L.bax_comm_destroy(comm)
            comm = None
        _native_comms[key] = comm
    return _native_comms[key]


class BAProblem:
    """One BA problem bound to one GPU.  Arguments are what reference ba.objfun receives
    (ba.py:113): K (C,3,3), sp2d (C,N*J,2) or (C,N,J,2), ss2d (C,N,J), sp3d (C,N,J3,3), ss3d (C,N,J3).

    Multi-GPU: give every rank its own contiguous block of frames and a torch.distributed
    `group`; the small packed buffers are all-reduced between the kernel phases (include/bax.h)."""

    def __init__(self, K, sp2d, ss2d, sp3d, ss3d, bone_idx=OP_BONE, lambda1=1.0, lambda2=1.0,
                 device=None, group=None, n_frames_global=None, handle=None):
        if not torch.cuda.is_available():
            raise BaxError("no CUDA device: this package has no CPU fallback")
        self.L = lib()
        self.dev = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        ss2d_t = torch.as_tensor(ss2d)
        C_, N_, J_ = ss2d_t.shape
        self.C, self.N, self.J = int(C_), int(N_), int(J_)
        self.P = self.N * self.J
        f = dict(dtype=torch.float64, device=self.dev)
        with torch.cuda.device(self.dev):
            self.K = torch.as_tensor(K).to(**f).contiguous()
            self.y = torch.as_tensor(sp2d).to(**f).contiguous().reshape(self.C, self.P, 2)
            self.s = ss2d_t.to(**f).contiguous().reshape(self.C, self.P)
            p3 = torch.as_tensor(sp3d).to(**f).contiguous() if sp3d is not None else None
            s3 = torch.as_tensor(ss3d).to(**f).contiguous() if ss3d is not None else None
            self.J3 = int(p3.shape[2]) if p3 is not None else self.J
            bones = np.ascontiguousarray(np.asarray(bone_idx), dtype=np.int32)
            self.B = int(bones.shape[0])
            self.bones = bones
            self.lambda1, self.lambda2 = float(lambda1), float(lambda2)
            if handle is not None:                 # a handle detach()ed from a finished problem (sweeps: ba_batch)
                h = handle
            else:
                h = C.c_void_p()
                check(self.L.bax_create(self.dev.index or 0, C.byref(h)), "bax_create")
            self.h = h
            check(self.L.bax_problem_set(self.h, self.C, self.N, self.J, self.J3, self.B,
                                         self.K.data_ptr(), self.y.data_ptr(), self.s.data_ptr(),
                                         p3.data_ptr() if p3 is not None else None,
                                         s3.data_ptr() if s3 is not None else None,
                                         bones.ctypes.data_as(_lib.P_i32), self.lambda1, self.lambda2, _stream()),
                  "bax_problem_set")
            d = (C.c_int64 * 6)()
            check(self.L.bax_problem_dims(self.h, d), "bax_problem_dims")
            self.n_vis, self.n_v3d, self.m, self.n, self.aug_rows, _ = [int(v) for v in d]
            self.step = torch.zeros(self.n, **f)
        self.group = group
        self.comm_events = None
        self._views = {}
        self.native_comm = None
        if group is not None:
            self.native_comm = _native_comm(group, self.dev)
            if self.native_comm is not None:
                check(self.L.bax_allreduce_attach(self.h, self.native_comm), "bax_allreduce_attach")
        if n_frames_global is not None:
            check(self.L.bax_set_global_frames(self.h, int(n_frames_global)), "bax_set_global_frames")
        self._out4 = (C.c_double * 8)()
        self.h_scale = 0.0
        self.n_trials = 0                     # damped solves so far (bench accounting)

    def rebind_observations(self, y, s):
        """Use another device copy of (sp2d, ss2d) — same shapes and visibility pattern — from now on."""
        assert y.is_cuda and s.is_cuda and y.dtype == torch.float64 and s.dtype == torch.float64
        assert y.numel() == self.y.numel() and s.numel() == self.s.numel() and y.is_contiguous() and s.is_contiguous()
        check(self.L.bax_observations_rebind(self.h, y.data_ptr(), s.data_ptr()), "bax_observations_rebind")
        self.y, self.s = y.reshape(self.C, self.P, 2), s.reshape(self.C, self.P)

    def detach(self):
        """Give up the libbax handle without destroying it: the next BAProblem(handle=...) on the same device reuses
        its pinned staging, events and stream-ordered memory (bax_problem_set releases the old problem itself)."""
        h, self.h = self.h, None
        return h

    def close(self):
        if getattr(self, "h", None):
            self.L.bax_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- helpers
    def _dev_theta(self, theta):
        if isinstance(theta, torch.Tensor) and theta.device == self.dev and theta.dtype == torch.float64:
            return theta.contiguous()
        return torch.as_tensor(np.asarray(theta, dtype=np.float64)).to(self.dev)

    def _allreduce(self, which, op=None):
        if self.group is None:
            return
        import torch.distributed as dist
        p, n = C.c_void_p(), C.c_int64()
        check(self.L.bax_reduce_buffer(self.h, which, C.byref(p), C.byref(n)), "bax_reduce_buffer")
        key = (which, p.value)                 # the LIN buffer changes address on linearize_push / _pop
        if key not in self._views:
            self._views[key] = torch.as_tensor(_DevView(p.value, n.value), device=self.dev)
        if self.comm_events is not None:            # bench instrumentation: device time spent in (and waiting for) the collective
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            dist.all_reduce(self._views[key], op=op or dist.ReduceOp.SUM, group=self.group)
            e1.record()
            self.comm_events.append((which, n.value, e0, e1))
            return
        dist.all_reduce(self._views[key], op=op or dist.ReduceOp.SUM, group=self.group)

    def comm_timing_read(self):
        """{buffer: (calls, total ms, doubles per call)} of the all-reduces recorded since comm_events was set to []."""
        torch.cuda.synchronize(self.dev)
        out = {}
        for which, cnt, e0, e1 in self.comm_events or []:
            c, ms, _ = out.get(which, (0, 0.0, cnt))
            out[which] = (c + 1, ms + e0.elapsed_time(e1), cnt)
        return out

    # ---- reference-shaped evaluations
    def residual(self, theta):
        """ba.objfun(theta, ...) -> (m,) torch tensor on the device."""
        th = self._dev_theta(theta)
        out = torch.empty(self.m, dtype=torch.float64, device=self.dev)
        with torch.cuda.device(self.dev):
            check(self.L.bax_residual(self.h, th.data_ptr(), out.data_ptr(), _stream()), "bax_residual")
        self._keep = th
        return out

    def residual_jacobian(self, theta, with_index=True, out=None):
        """Fused residual + Jacobian blocks of the visible reprojection rows (device tensors):
        r2 (n_vis,2), Jcam (n_vis,2,6), Jpt (n_vis,2,3), obs_cam, obs_pt (n_vis,) int32."""
        th = self._dev_theta(theta)
        f = dict(dtype=torch.float64, device=self.dev)
        if out is None:
            r2 = torch.empty(self.n_vis, 2, **f)
            Jc = torch.empty(self.n_vis, 2, 6, **f)
            Jp = torch.empty(self.n_vis, 2, 3, **f)
            oc = torch.empty(self.n_vis, dtype=torch.int32, device=self.dev) if with_index else None
            op = torch.empty(self.n_vis, dtype=torch.int32, device=self.dev) if with_index else None
        else:
            r2, Jc, Jp, oc, op = out
        with torch.cuda.device(self.dev):
            check(self.L.bax_residual_jacobian(self.h, th.data_ptr(), r2.data_ptr(), Jc.data_ptr(), Jp.data_ptr(),
                                               oc.data_ptr() if oc is not None else None,
                                               op.data_ptr() if op is not None else None, _stream()),
                  "bax_residual_jacobian")
        self._keep = th
        return r2, Jc, Jp, oc, op

    def jacobian(self, theta):
        """Analytic Jacobian of ba.objfun as scipy CSR (m x n), assembled on the host from the device
        blocks.  (The reference has none: SciPy finite-differences ba.objfun, ba.py:187-206.)"""
        import scipy.sparse as sp
        th = self._dev_theta(theta)
        r2, Jc, Jp, oc, op = self.residual_jacobian(th)
        Cn, nv = self.C, self.n_vis
        oc = oc.cpu().numpy().astype(np.int64); op = op.cpu().numpy().astype(np.int64)
        Jc = Jc.cpu().numpy(); Jp = Jp.cpu().numpy()
        rows = (2 * np.arange(nv)[:, None] + np.arange(2)[None, :])              # (nv,2)
        cols_r = 3 * oc[:, None] + np.arange(3)[None, :]
        cols_t = 3 * Cn + 3 * oc[:, None] + np.arange(3)[None, :]
        cols_p = 6 * Cn + 3 * op[:, None] + np.arange(3)[None, :]
        R_, C_, V_ = [], [], []
        for cols, blk in ((cols_r, Jc[:, :, :3]), (cols_t, Jc[:, :, 3:]), (cols_p, Jp)):
            R_.append(np.broadcast_to(rows[:, :, None], (nv, 2, 3)).reshape(-1))
            C_.append(np.broadcast_to(cols[:, None, :], (nv, 2, 3)).reshape(-1))
            V_.append(blk.reshape(-1))
        f = dict(dtype=torch.float64, device=self.dev)
        Jv = torch.zeros(max(self.n_v3d, 1), 3 * Cn, **f)
        Jb = torch.zeros(self.N, max(self.B, 1), 3, **f)
        with torch.cuda.device(self.dev):
            check(self.L.bax_prior_jacobian(self.h, th.data_ptr(), Jv.data_ptr(), Jb.data_ptr(), _stream()),
                  "bax_prior_jacobian")
        if self.n_v3d:
            Jv = Jv.cpu().numpy()
            R_.append(np.repeat(2 * nv + np.arange(self.n_v3d), 3 * Cn))
            C_.append(np.tile(np.arange(3 * Cn), self.n_v3d))
            V_.append(Jv.reshape(-1))
        if self.B:
            Jb = Jb.cpu().numpy()
            row0 = 2 * nv + self.n_v3d
            n_idx = np.arange(self.N)
            for b in range(self.B):
                for e, sg in ((0, 1.0), (1, -1.0)):
                    pcol = 6 * Cn + 3 * (n_idx * self.J + int(self.bones[b, e]))
                    R_.append(np.full(3 * self.N, row0 + b))
                    C_.append((pcol[:, None] + np.arange(3)[None, :]).reshape(-1))
                    V_.append((sg * Jb[:, b, :]).reshape(-1))
        return sp.csr_matrix((np.concatenate(V_), (np.concatenate(R_), np.concatenate(C_))),
                             shape=(self.m, self.n))

    # ---- LM building blocks (phases split where a frame-sharded run all-reduces)
    def cost(self, theta):
        th = self._dev_theta(theta)
        self._keep_cost = th
        st = _stream()
        with torch.cuda.device(self.dev):
            check(self.L.bax_cost_begin(self.h, th.data_ptr(), st), "bax_cost_begin")
            self._allreduce(_lib.BUF_BONE)
            check(self.L.bax_cost_mid(self.h, st), "bax_cost_mid")
            self._allreduce(_lib.BUF_COST)
            out = C.c_double()
            check(self.L.bax_cost_finish(self.h, C.byref(out), st), "bax_cost_finish")
        return out.value

    def linearize(self, theta):
        """-> (cost, ||g||_inf, ||g||_2, ||theta||_2); keeps V_p, g_p, H_cc on the device."""
        th = self._dev_theta(theta)
        self._keep_lin = th
        st = _stream()
        with torch.cuda.device(self.dev):
            check(self.L.bax_linearize_begin(self.h, th.data_ptr(), st), "bax_linearize_begin")
            self._allreduce(_lib.BUF_BONE)
            check(self.L.bax_linearize_mid(self.h, st), "bax_linearize_mid")
            self._allreduce(_lib.BUF_LIN)
            check(self.L.bax_linearize_post(self.h, st), "bax_linearize_post")
            if self.group is not None:
                import torch.distributed as dist
                self._allreduce(_lib.BUF_NORM)
                self._allreduce(_lib.BUF_GMAX, dist.ReduceOp.MAX)
            check(self.L.bax_linearize_finish(self.h, self._out4, st), "bax_linearize_finish")
        self.h_scale = self._out4[4]          # max diagonal of the point blocks of J^T J
        return tuple(self._out4[:4])

    def linearize_push(self):
        """Save the current linearisation; the next linearize() may run at a trial point."""
        check(self.L.bax_linearize_push(self.h), "bax_linearize_push")
        self._saved_lin = (getattr(self, "_keep_lin", None), self.h_scale)

    def linearize_pop(self):
        """Restore the linearisation saved by linearize_push() (the trial step was rejected)."""
        check(self.L.bax_linearize_pop(self.h), "bax_linearize_pop")
        self._keep_lin, self.h_scale = self._saved_lin

    def trial(self, alpha, need_phi_prime=True):
        """Solve (J^T J + alpha I) p = -g; p lands in self.step (device).
        -> (||p||^2, g.p, p^T (J^T J + alpha I)^-1 p)."""
        st = _stream()
        self.n_trials += 1
        with torch.cuda.device(self.dev):
            check(self.L.bax_trial_begin(self.h, float(alpha), st), "bax_trial_begin")
            self._allreduce(_lib.BUF_SCHUR)
            check(self.L.bax_trial_mid(self.h, self.step.data_ptr(), 1 if need_phi_prime else 0, st), "bax_trial_mid")
            self._allreduce(_lib.BUF_RHS2)
            check(self.L.bax_trial_finish(self.h, self._out4, st), "bax_trial_finish")
        pn2, gp, ptq = self._out4[0], self._out4[1], self._out4[2]
        if not (math.isfinite(pn2) and math.isfinite(gp)):
            raise BaxError(f"non-finite trial step at alpha={alpha}: reduced camera system not SPD")
        return pn2, gp, ptq

    def trial_phi_prime(self):
        """p^T (J^T J + alpha I)^-1 p of the last trial(alpha, False), computed on demand (second right-hand side from
        the whitened blocks still on the device + second reduced solve)."""
        st = _stream()
        with torch.cuda.device(self.dev):
            check(self.L.bax_trial_phi_prime(self.h, st), "bax_trial_phi_prime")
            self._allreduce(_lib.BUF_RHS2_HEAD)
            check(self.L.bax_trial_finish(self.h, self._out4, st), "bax_trial_finish")
        return self._out4[2]

    def triangulate(self, R_w2c, t_w2c):
        """util.triangulate_with_conf (reference util.py:325-354) -> (N,J,3) numpy."""
        f = dict(dtype=torch.float64, device=self.dev)
        R = torch.as_tensor(np.asarray(R_w2c, dtype=np.float64)).to(**f).contiguous()
        t = torch.as_tensor(np.asarray(t_w2c, dtype=np.float64).reshape(self.C, 3)).to(**f).contiguous()
        X = torch.empty(self.P, 3, **f)
        st = C.c_int32(0)
        with torch.cuda.device(self.dev):
            check(self.L.bax_triangulate(self.h, R.data_ptr(), t.data_ptr(), X.data_ptr(), C.byref(st), _stream()),
                  "bax_triangulate")
        if st.value == 1:
            raise Exception("not enough constraint")          # reference util.py:320-321
        return X.cpu().numpy().reshape(self.N, self.J, 3)

    def reproject_error(self, R_w2c, t_w2c, X):
        """Pixel distance between every observation and the projection of its point (eval.py:76-108, ba.py:248-256)
        -> (C, N, J) numpy, NaN where invisible."""
        f = dict(dtype=torch.float64, device=self.dev)
        R = torch.as_tensor(np.asarray(R_w2c, dtype=np.float64)).to(**f).contiguous()
        t = torch.as_tensor(np.asarray(t_w2c, dtype=np.float64).reshape(self.C, 3)).to(**f).contiguous()
        Xd = torch.as_tensor(np.asarray(X, dtype=np.float64).reshape(self.P, 3)).to(**f).contiguous()
        e = torch.empty(self.C, self.P, **f)
        with torch.cuda.device(self.dev):
            check(self.L.bax_reproject_error(self.h, R.data_ptr(), t.data_ptr(), Xd.data_ptr(), e.data_ptr(), _stream()),
                  "bax_reproject_error")
        return e.cpu().numpy().reshape(self.C, self.N, self.J)

    # ---- timing hooks
    def timing(self, on=True):
        check(self.L.bax_timing_enable(self.h, 1 if on else 0))
        check(self.L.bax_timing_reset(self.h))

    def timing_read(self):
        ms = (C.c_double * 8)(); nl = (C.c_int64 * 8)()
        check(self.L.bax_timing_read(self.h, ms, nl))
        return {k: (ms[i], int(nl[i])) for i, k in enumerate(_lib.T_NAMES)}


# ----------------------------------------------------------------------------- trust-region driver
def _solve_tr(prob, Delta, g_norm2, initial_alpha, rtol=0.01, max_iter=10):
    """More'-Sorensen root-finding of ||p(alpha)|| = Delta (scipy/optimize/_lsq/common.py:57-168), with
    every phi / phi' evaluation a damped Schur solve on the GPU instead of an SVD of the dense Jacobian.

    The exact Jacobian has a 6-dimensional gauge null space (plus a nearly flat global-scale mode), so
    SciPy's `full_rank` branch never applies: alpha_lower starts at 0.  When the minimum-norm Gauss-Newton
    step lies inside the region (phi < 0 however small alpha gets) SciPy's loop drives alpha to round-off
    level and follows noise in the null-space singular values; here alpha is floored at
    1e-13 x the Hessian scale and that interior step is returned unscaled.
    -> (alpha, iterations, ||p||, g.p, on_boundary)"""
    alpha_floor = 1e-13 * prob.h_scale
    alpha_upper = g_norm2 / Delta
    alpha_lower = 0.0

    def reset():
        return max(0.001 * alpha_upper, (alpha_lower * alpha_upper) ** 0.5, alpha_floor)

    if initial_alpha is None or initial_alpha == 0:
        # first solve of a run: Delta_0 = ||x_0|| is far larger than any sensible step, so probe the floor first —
        # if that (Gauss-Newton-like) step is interior, the loop below would only walk alpha down to it
        pn2, gp, _ = prob.trial(alpha_floor, False)
        if math.sqrt(pn2) <= Delta:
            return alpha_floor, 0, math.sqrt(pn2), gp, False
        alpha = reset()
    else:
        alpha = max(initial_alpha, alpha_floor)
    it = 0
    for it in range(max_iter):
        if alpha < alpha_lower or alpha > alpha_upper:
            alpha = reset()
        alpha = max(alpha, alpha_floor)
        # phi' costs another pass over the whitened blocks and a second reduced solve: at the floor the step is usually
        # the interior one and returns before phi' is looked at, so it is computed only once that is ruled out
        lazy = alpha <= alpha_floor
        pn2, gp, ptq = prob.trial(alpha, not lazy)
        p_norm = math.sqrt(pn2)
        phi = p_norm - Delta
        if phi < 0 and alpha <= alpha_floor:
            return alpha, it + 1, p_norm, gp, False          # interior (Gauss-Newton-like) step
        if lazy:
            ptq = prob.trial_phi_prime()
        if not (ptq > 0 and p_norm > 0):
            break                                             # step underflowed (start is already stationary)
        phi_prime = -ptq / p_norm
        if phi < 0:
            alpha_upper = alpha
        ratio = phi / phi_prime
        alpha_lower = max(alpha_lower, alpha - ratio)
        alpha -= (phi + Delta) * ratio / Delta
        if abs(phi) < rtol * Delta:
            break
    alpha = max(alpha, alpha_floor)
    pn2, gp, _ = prob.trial(alpha, False)
    return alpha, it + 1, math.sqrt(pn2), gp, True


def _update_tr_radius(Delta, actual, predicted, step_norm, bound_hit):
    """scipy/optimize/_lsq/common.py:222-248."""
    if predicted > 0:
        ratio = actual / predicted
    elif predicted == actual == 0:
        ratio = 1
    else:
        ratio = 0
    if ratio < 0.25:
        Delta = 0.25 * step_norm
    elif ratio > 0.75 and bound_hit:
        Delta *= 2.0
    return Delta, ratio


def _check_termination(dF, F, dx_norm, x_norm, ratio, ftol, xtol):
    """scipy/optimize/_lsq/common.py:705-717."""
    ftol_ok = dF < ftol * F and ratio > 0.25
    xtol_ok = dx_norm < xtol * (xtol + x_norm)
    if ftol_ok and xtol_ok:
        return 4
    if ftol_ok:
        return 2
    if xtol_ok:
        return 3
    return None


TERMINATION_MESSAGES = {0: "The maximum number of function evaluations is exceeded.",
                        1: "`gtol` termination condition is satisfied.",
                        2: "`ftol` termination condition is satisfied.",
                        3: "`xtol` termination condition is satisfied.",
                        4: "Both `ftol` and `xtol` termination conditions are satisfied."}


class TRFSolver:
    """Trust-region-reflective loop of SciPy's `trf_no_bounds` (scipy/optimize/_lsq/trf.py:415-587,
    x_scale = 1, tr_solver = 'exact') — the solver reference ba.py:187-206 runs — with the Jacobian,
    normal equations and damped solves on the GPU.  `iterate()` is one pass of SciPy's outer loop body:
    trust-region solves and cost evaluations until a step is accepted (or evaluations run out), then a
    re-linearisation."""

    def __init__(self, prob, theta0, ftol=1e-4, xtol=1e-8, gtol=1e-8, max_nfev=None, speculative=True, native=None):
        self.prob = prob
        # speculative: a trial point is linearised straight away (its cost falls out of the linearisation), so an
        # accepted step needs no separate cost pass; a rejected step restores the saved linearisation
        self.speculative = speculative
        # native: the same loop behind one C-ABI call per outer iteration (csrc/bax_trf.cu); a frame-sharded run
        # all-reduces between the phases inside the library too, through the communicator attached to the handle
        if native is None:
            native = os.environ.get("BAX_TRF_HOST", "0") != "1"
        # (a frame-sharded run needs libbax's own communicator for that; an instrumented pass — comm_events — stays here)
        self.native = bool(native) and speculative and (prob.group is None or
                                                        (prob.native_comm is not None and prob.comm_events is None))
        self.x = prob._dev_theta(theta0).clone()
        self.x_new = torch.empty_like(self.x)
        self.ftol, self.xtol, self.gtol = ftol, xtol, gtol
        self.max_nfev = 100 * prob.n if max_nfev is None else max_nfev
        self.history = []
        if self.native:
            st = self._st = _lib.TrfState()
            st.ftol, st.xtol, st.gtol, st.max_nfev = ftol, xtol, gtol, int(self.max_nfev)
            st.x, st.x_new, st.step = self.x.data_ptr(), self.x_new.data_ptr(), prob.step.data_ptr()
            with torch.cuda.device(prob.dev):
                rc = prob.L.bax_trf_begin(prob.h, C.byref(st), _stream())
            if rc != 0 and not math.isfinite(st.cost):
                raise ValueError("Residuals are not finite in the initial point.")
            check(rc, "bax_trf_begin")
            self._pull()
            return
        self.cost, self.g_inf, self.g_2, self.x_norm = prob.linearize(self.x)
        if not math.isfinite(self.cost):
            raise ValueError("Residuals are not finite in the initial point.")
        self.initial_cost = self.cost
        self.nfev, self.njev = 1, 1
        self.Delta = self.x_norm if self.x_norm > 0 else 1.0
        self.alpha = 0.0
        self.status = None
        self.iteration = 0

    # ---- native state <-> attributes (the attributes stay the interface: bench.py and the tests set Delta / alpha)
    def _pull(self):
        st = self._st
        self.cost, self.g_inf, self.g_2, self.x_norm = st.cost, st.g_inf, st.g_2, st.x_norm
        self.initial_cost, self.Delta, self.alpha = st.initial_cost, st.Delta, st.alpha
        self.nfev, self.njev, self.iteration = int(st.nfev), int(st.njev), int(st.iteration)
        self.status = None if st.status < 0 else int(st.status)
        self.prob.h_scale = st.h_scale
        if st.x != self.x.data_ptr():                      # an accepted step swapped the roles of the two vectors
            self.x, self.x_new = self.x_new, self.x
        self.prob._keep_lin = self.x

    def _push(self):
        st = self._st
        st.cost, st.g_inf, st.g_2, st.x_norm, st.h_scale = self.cost, self.g_inf, self.g_2, self.x_norm, self.prob.h_scale
        st.Delta, st.alpha = self.Delta, self.alpha
        st.nfev, st.njev, st.iteration = self.nfev, self.njev, self.iteration
        st.status = -1 if self.status is None else self.status
        st.x, st.x_new = self.x.data_ptr(), self.x_new.data_ptr()

    def _iterate_native(self):
        st, prob = self._st, self.prob
        self._push()
        n0 = int(st.n_trials)
        with torch.cuda.device(prob.dev):
            check(prob.L.bax_trf_iterate(prob.h, C.byref(st), _stream()), "bax_trf_iterate")
        prob.n_trials += int(st.n_trials) - n0
        for k in range(st.n_hist):
            r = st.hist[k]
            self.history.append(dict(alpha=r[0], Delta=r[1], step_norm=r[2], cost_new=r[3], pred=r[4], act=r[5], n_it=int(r[6])))
        self._pull()

    def done(self):
        if self.status is None and self.g_inf < self.gtol:
            self.status = 1
        return self.status is not None or self.nfev >= self.max_nfev

    def iterate(self, relinearize=True):
        if self.native and relinearize:
            return self._iterate_native()
        prob = self.prob
        actual_reduction = -1.0
        cost_new = self.cost
        while actual_reduction <= 0 and self.nfev < self.max_nfev:
            self.alpha, n_it, p_norm, gdotp, on_boundary = _solve_tr(prob, self.Delta, self.g_2, self.alpha)
            alpha = self.alpha
            kappa = self.Delta / p_norm if on_boundary else 1.0   # SciPy rescales a boundary step onto the boundary
            # -(0.5 ||J s||^2 + g.s) with s = kappa p and p^T J^T J p = -g.p - alpha ||p||^2
            predicted = -(0.5 * kappa * kappa * (-gdotp - alpha * p_norm * p_norm) + kappa * gdotp)
            torch.add(self.x, prob.step, alpha=kappa, out=self.x_new)
            spec = self.speculative and relinearize
            if spec:
                prob.linearize_push()
                cost_new, g_inf_new, g_2_new, x_norm_new = prob.linearize(self.x_new)
            else:
                cost_new = prob.cost(self.x_new)
            self.nfev += 1
            step_norm = kappa * p_norm
            if not math.isfinite(cost_new):
                if spec:
                    prob.linearize_pop()
                self.Delta = 0.25 * step_norm
                continue
            actual_reduction = self.cost - cost_new
            if spec and not actual_reduction > 0:
                prob.linearize_pop()
            Delta_new, ratio = _update_tr_radius(self.Delta, actual_reduction, predicted, step_norm,
                                                 step_norm > 0.95 * self.Delta)
            self.history.append(dict(alpha=alpha, Delta=self.Delta, step_norm=step_norm, cost_new=cost_new,
                                     pred=predicted, act=actual_reduction, n_it=n_it))
            self.status = _check_termination(actual_reduction, self.cost, step_norm, self.x_norm, ratio,
                                             self.ftol, self.xtol)
            if self.status is not None:
                break
            if not on_boundary:
                # SciPy returns alpha = 0 for an interior Gauss-Newton step (common.py:113-116) and starts the next
                # sub-problem from there; carrying the floor value instead costs a wasted damped solve whenever the
                # Hessian scale (hence the floor) drops a little between linearisations
                self.alpha = 0.0
            self.alpha *= self.Delta / Delta_new
            self.Delta = Delta_new
        if actual_reduction > 0:
            self.x, self.x_new = self.x_new, self.x
            if spec:
                self.cost, self.g_inf, self.g_2, self.x_norm = cost_new, g_inf_new, g_2_new, x_norm_new
                self.njev += 1
            elif relinearize:
                self.cost, self.g_inf, self.g_2, self.x_norm = prob.linearize(self.x)
                self.njev += 1
            else:
                self.cost = cost_new
        self.iteration += 1

    def solve_native(self):
        """All remaining iterations behind ONE call across the ABI (bax_trf_iterate until termination, in C)."""
        st, prob = self._st, self.prob
        self._push()
        n0 = int(st.n_trials)
        with torch.cuda.device(prob.dev):
            check(prob.L.bax_trf_solve(prob.h, C.byref(st), _stream()), "bax_trf_solve")
        prob.n_trials += int(st.n_trials) - n0
        self._pull()

    def result(self):
        status = 0 if self.status is None else self.status
        return dict(x=self.x, cost=self.cost, optimality=self.g_inf, nfev=self.nfev, njev=self.njev,
                    status=status, iterations=self.iteration, history=self.history,
                    message=TERMINATION_MESSAGES[status], initial_cost=self.initial_cost)


def trf_schur(prob, theta0, ftol=1e-4, xtol=1e-8, gtol=1e-8, max_nfev=None, verbose=0, callback=None):
    """Run TRFSolver to termination.  Returns a dict like SciPy's OptimizeResult (x stays on the device)."""
    sol = TRFSolver(prob, theta0, ftol, xtol, gtol, max_nfev)
    if sol.native and callback is None:
        sol.solve_native()
    while not sol.done():
        sol.iterate()
        if callback is not None:
            callback(sol.iteration, sol.cost, sol.nfev)
    res = sol.result()
    if verbose:
        print(res["message"])
        print(f"Function evaluations {res['nfev']}, initial cost {res['initial_cost']:.4e}, "
              f"final cost {res['cost']:.4e}, first-order optimality {res['optimality']:.2e}.")
    return res


# ----------------------------------------------------------------------------- reference entry points
_cache = {}


def _fingerprint(a):
    """Content hash of an array: the drop-ins below are called with the same constants thousands of times (SciPy's
    finite differences), but a caller may also mutate an array in place between calls (the reference's own
    joints2projections does, util.py:556-569) — identity of the object says nothing about its content."""
    import hashlib
    if a is None:
        return None
    a = np.ascontiguousarray(a)
    return (a.shape, a.dtype.str, hashlib.blake2b(a.view(np.uint8).reshape(-1), digest_size=16).digest())


def _problem_for(K, sp2d, ss2d, sp3d, ss3d, bone_idx, C, N, J, lambda1, lambda2):
    key = (_fingerprint(K), _fingerprint(sp2d), _fingerprint(ss2d), _fingerprint(sp3d), _fingerprint(ss3d),
           _fingerprint(np.asarray(bone_idx)), C, N, J, float(lambda1), float(lambda2))
    hit = _cache.get("p")
    if hit is not None and hit[0] == key:
        return hit[1]
    if hit is not None:
        hit[1].close()
    prob = BAProblem(K, np.asarray(sp2d).reshape(C, N * J, 2), np.asarray(ss2d).reshape(C, N, J), sp3d, ss3d,
                     bone_idx, lambda1, lambda2)
    _cache["p"] = (key, prob)
    return prob


def objfun(params, K, sp2d, ss2d, sp3d, ss3d, bone_idx, C, N, J, lambda1, lambda2):
    """Drop-in for reference ba.objfun (ba.py:113-140): NumPy in, NumPy (m,) out, computed on the GPU."""
    prob = _problem_for(K, sp2d, ss2d, sp3d, ss3d, bone_idx, C, N, J, lambda1, lambda2)
    return prob.residual(params).cpu().numpy()


def jacobian(params, K, sp2d, ss2d, sp3d, ss3d, bone_idx, C, N, J, lambda1, lambda2):
    """Analytic Jacobian matching `objfun` (scipy CSR); usable as least_squares(jac=...)."""
    prob = _problem_for(K, sp2d, ss2d, sp3d, ss3d, bone_idx, C, N, J, lambda1, lambda2)
    return prob.jacobian(params)


def ba_main(camid, K, R_w2c, t_w2c, sp2d, ss2d, sp3d, ss3d, lambda1, lambda2, *,
            ftol=1e-4, xtol=1e-8, gtol=1e-8, max_nfev=None, verbose=True, return_info=False, _reuse=None):
    """Drop-in for reference ba.ba_main (ba.py:157-210): triangulate the start, print the three initial
    mean costs, minimise ba.objfun by trust-region least squares (ftol=1e-4), unpack."""
    C = len(camid)
    N, J = sp2d.shape[1], sp2d.shape[2]
    prob = BAProblem(K, np.asarray(sp2d).reshape(C, N * J, 2), ss2d, sp3d, ss3d, OP_BONE, lambda1, lambda2,
                     handle=_reuse.pop("h", None) if _reuse is not None else None)
    try:
        x_all = prob.triangulate(R_w2c, t_w2c).reshape(N * J, 3)
        theta0 = to_theta(R_w2c, t_w2c, x_all)
        if not np.all(np.isfinite(theta0)):
            # NaN from an unseen / single-view point reaches SciPy's bounds check in the reference
            raise ValueError("Initial guess is outside of provided bounds")
        if verbose:
            f0 = prob.residual(theta0).cpu().numpy()
            e1 = f0[:2 * prob.n_vis]
            if lambda1 == 0 or lambda2 == 0:
                # the reference prints the UNSCALED prior terms (ba.py:179-183); with a zero weight they cannot be
                # recovered from the weighted residual, so evaluate them once with unit weights
                unit = BAProblem(K, np.asarray(sp2d).reshape(C, N * J, 2), ss2d, sp3d, ss3d, OP_BONE, 1.0, 1.0)
                try:
                    f1 = unit.residual(theta0).cpu().numpy()
                finally:
                    unit.close()
                e2 = f1[2 * prob.n_vis:2 * prob.n_vis + prob.n_v3d]
                e3 = f1[2 * prob.n_vis + prob.n_v3d:]
            else:
                e2 = f0[2 * prob.n_vis:2 * prob.n_vis + prob.n_v3d] / lambda1
                e3 = f0[2 * prob.n_vis + prob.n_v3d:] / lambda2
            print(f"Initial mean NLL: {np.mean(e1 * e1)}")
            print(f"Initial mean 3D variance: {np.mean(e2 * e2)}")
            print(f"Initial mean bone-length variance: {np.mean(e3 * e3)}")
        res = trf_schur(prob, theta0, ftol=ftol, xtol=xtol, gtol=gtol, max_nfev=max_nfev, verbose=1 if verbose else 0)
        theta = res["x"].cpu().numpy()
        res["x"] = theta
    finally:
        if _reuse is not None:
            _reuse["h"] = prob.detach()         # the sweep's worker keeps the handle for its next problem
        else:
            prob.close()
    R, t, x = from_theta(theta, C)
    return (R, t, x, res) if return_info else (R, t, x)


# ----------------------------------------------------------------------------- batches of independent problems
def batch_plan(n_problems, rank=0, world=1):
    """Problem indices owned by `rank`: round-robin (problem k -> rank k % world), so that a sweep whose cost
    drifts with the index (longer sequences last) still loads every GPU alike.  Replicas only — the problems
    of BASELINE configs[4] (4096 independent 4-camera scenes) share nothing, so there is no collective on the
    data path; only the results are gathered."""
    return list(range(rank, n_problems, world))


_worker_streams = {}


def _worker_stream(dev, i):
    """The i-th worker stream of device `dev`, created once per process: torch's caching allocator keeps its blocks
    per stream, so a sweep that came back with fresh streams would pay cudaMalloc for its uploads again."""
    pool = _worker_streams.setdefault(dev, [])
    while len(pool) <= i:
        pool.append(torch.cuda.Stream(device=dev))
    return pool[i]


batch_stats = {}          # what the last native sweep of this process used: workers per GPU, wait mode, cores seen


def _ba_batch_native(idx, problems, lambda1, lambda2, n_workers, ftol=1e-4, xtol=1e-8, gtol=1e-8, max_nfev=None,
                     return_info=False):
    """The sweep behind ONE C call per problem shape (bax_batch_solve, csrc/bax_batch.cu): native worker threads, each
    with its own stream and libbax handle, run per problem the very sequence ba_main runs.  The interpreter only
    prepares pointers and converts all rotations at once (rodrigues_inv_batch / rodrigues_batch) — per-problem Python
    under the GIL (~1 ms) was what capped the threaded sweep at ~800 problems/s.  -> {index: result}."""
    L = lib()
    dev = torch.cuda.current_device()
    f8 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    groups = {}
    for k in idx:
        p = problems[k]
        ss2d = np.asarray(p[5])
        has3 = p[6] is not None and p[7] is not None
        key = (len(p[0]),) + tuple(ss2d.shape) + ((np.asarray(p[7]).shape[2],) if has3 else (0,))
        groups.setdefault(key, []).append(k)
    out = {}
    bones = np.ascontiguousarray(OP_BONE, dtype=np.int32)
    # worker threads spin on their scalar read-backs: one per core this rank can have.  With several ranks on one
    # host (LOCAL_WORLD_SIZE) the cores are shared: 4 ranks x 12 spinning workers on a small host ran at a third of
    # the single-rank rate.  Measured with 4 cores per rank: 3 workers 957, 4 workers 1163, 8 workers 1099 problems/s
    # per rank spinning; sleeping waits (bax_set_blocking_sync) 627 / 744 / 996 — so: spin, workers = cores per rank
    # (at least 4), sleep only when even that oversubscribes the host twice over.
    try:
        cpus = len(os.sched_getaffinity(0))
    except AttributeError:
        cpus = os.cpu_count() or 1
    local_world = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    nw = max(1, min(int(n_workers), max(4, cpus // local_world)))
    blocking = 1 if nw * local_world > 2 * cpus else 0
    if os.environ.get("BAX_BATCH_FORCE_WORKERS"):             # experiments: fixed worker count / wait mode
        nw = int(os.environ["BAX_BATCH_FORCE_WORKERS"])
        blocking = 1 if os.environ.get("BAX_BATCH_BLOCK", "0") == "1" else 0
    batch_stats.update(workers=nw, blocking_sync=int(blocking), cpus=cpus, local_world=local_world)
    # The interpreter's share (contiguity checks, pointer tables, the Rodrigues conversions, result tuples: ~0.37 ms per
    # problem) is serial; the C call releases the GIL.  So the problems of a shape go through in chunks: while the
    # native workers solve chunk i, this thread prepares chunk i + 1 and unpacks chunk i - 1 (measured on 512 problems,
    # 12 workers: 0.45 s -> see DESIGN.md 8b).  BAX_BATCH_CHUNK=0 restores one call per shape.
    chunk = int(os.environ.get("BAX_BATCH_CHUNK", str(max(64, 8 * nw))))
    from concurrent.futures import ThreadPoolExecutor
    for (Cn, C2, N, J, J3), ks_all in groups.items():
        assert Cn == C2
        keep = {}

        def arr(a, shape):                      # one contiguous float64 copy per distinct array object
            key = id(a)
            if key not in keep:
                b = f8(a).reshape(shape)
                keep[key] = (a, b)
            return keep[key][1]

        def prep(ks):
            n = len(ks)
            cols = {name: [] for name in ("K", "y", "s", "p3", "s3", "R", "t")}
            Rall = np.empty((n, Cn, 3, 3))
            for i, k in enumerate(ks):
                camid, K, R, t, sp2d, ss2d, sp3d, ss3d = problems[k]
                cols["K"].append(arr(K, (Cn, 3, 3))); cols["y"].append(arr(sp2d, (Cn, N * J, 2)))
                cols["s"].append(arr(ss2d, (Cn, N * J)))
                if J3:
                    cols["p3"].append(arr(sp3d, (Cn, N, J3, 3))); cols["s3"].append(arr(ss3d, (Cn, N, J3)))
                cols["R"].append(arr(R, (Cn, 3, 3))); cols["t"].append(arr(t, (Cn, 3)))
                Rall[i] = cols["R"][-1]
            rvec = np.ascontiguousarray(rodrigues_inv_batch(Rall.reshape(-1, 3, 3)).reshape(n, Cn, 3))
            ptrs = lambda xs: (C.c_void_p * n)(*[x.ctypes.data for x in xs])
            theta = np.empty((n, 6 * Cn + 3 * N * J))
            info = np.zeros((n, 8))
            a = _lib.BatchArgs()
            a.blocking_sync = blocking
            a.n_problems, a.n_workers, a.device = n, nw, dev
            a.C, a.N, a.J, a.J3, a.B = Cn, N, J, J3 or J, int(bones.shape[0])
            live = [cols, rvec, theta, info, ptrs(cols["K"]), ptrs(cols["y"]), ptrs(cols["s"]), ptrs(cols["R"]),
                    ptrs(cols["t"]), (C.c_void_p * n)(*[rvec[i].ctypes.data for i in range(n)])]
            a.K, a.y2d, a.s2d, a.R0, a.t0, a.rvec0 = live[4:10]
            if J3:
                live += [ptrs(cols["p3"]), ptrs(cols["s3"])]
                a.p3d, a.s3d = live[10], live[11]
            a.bones = bones.ctypes.data_as(_lib.P_i32)
            a.lambda1, a.lambda2 = float(lambda1), float(lambda2)
            a.ftol, a.xtol, a.gtol = float(ftol), float(xtol), float(gtol)
            a.max_nfev = 0 if max_nfev is None else int(max_nfev)
            a.theta_out, a.info_out = theta.ctypes.data, info.ctypes.data
            return a, live, theta, info

        def solve(job):                         # runs on the helper thread: the library keeps its error text per thread
            torch.cuda.set_device(dev)
            rc = L.bax_batch_solve(C.byref(job[0]))
            msg = L.bax_last_error() if rc else None
            return rc, (msg.decode() if msg else "")

        def done(fut):
            rc, msg = fut.result()
            if rc != 0:
                raise BaxError(f"bax_batch_solve failed ({rc}): {msg}")

        def post(ks, job):
            _, _, theta, info = job
            Rs = rodrigues_batch(theta[:, :3 * Cn].reshape(-1, 3)).reshape(len(ks), Cn, 3, 3)
            for i, k in enumerate(ks):
                st = int(info[i, 0])
                if st == -10:
                    raise Exception(f"ba_batch: problem {k}: not enough constraint")          # reference util.py:320-321
                if st == -11:
                    raise ValueError(f"ba_batch: problem {k}: Initial guess is outside of provided bounds")
                th = theta[i]
                R, t, x = Rs[i].copy(), th[3 * Cn:6 * Cn].reshape(Cn, 3, 1).copy(), th[6 * Cn:].reshape(-1, 3).copy()
                if return_info:
                    res = dict(x=th.copy(), cost=info[i, 1], optimality=info[i, 6], nfev=int(info[i, 3]),
                               njev=int(info[i, 4]), status=st, iterations=int(info[i, 5]), history=[],
                               message=TERMINATION_MESSAGES[st], initial_cost=info[i, 2])
                    out[k] = (R, t, x, res)
                else:
                    out[k] = (R, t, x)

        parts = [ks_all[i:i + chunk] for i in range(0, len(ks_all), chunk)] if chunk > 0 else [ks_all]
        with ThreadPoolExecutor(max_workers=1) as pool:
            pending = None                       # (ks, job, future) of the chunk the native workers are on
            for ks in parts:
                job = prep(ks)
                if pending is not None:
                    done(pending[2])
                fut = pool.submit(solve, job)
                if pending is not None:
                    post(pending[0], pending[1])
                pending = (ks, job, fut)
            done(pending[2])
            post(pending[0], pending[1])
    return out


def ba_batch(problems, lambda1, lambda2, *, n_streams=12, group=None, gather=True, solve=None, **kw):
    """Multi-sequence sweep: `ba_main` on every entry of `problems` — a sequence of
    (camid, K, R_w2c, t_w2c, sp2d, ss2d, sp3d, ss3d) tuples (or callables returning one, so a sweep need not
    hold every scene in host memory) — each by the same code path and therefore with the same result as a
    single `ba_main` call.

    One GPU: `n_streams` worker threads, each with its own (persistent) CUDA stream and its own libbax handle, pull problem
    indices from a shared queue.  A 4-camera problem is launch- and latency-bound on its own (a few hundred CTAs
    per kernel, ~40 launches and 2-3 scalar read-backs per trust-region iteration); the streams overlap the
    kernels of one problem with the read-backs and host logic of the others.  ctypes releases the GIL for the
    duration of every libbax call, read-backs included.
    Several GPUs: pass a torch.distributed `group`; problems are dealt round-robin over the ranks (batch_plan),
    every rank solves its own, and with gather=True the results are all-gathered so that every rank returns
    the full list (index order).  -> list of (R, t, x) (or (R, t, x, info) with return_info=True)."""
    import queue
    import threading
    n = len(problems)
    rank, world = 0, 1
    if group is not None:
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
    mine = batch_plan(n, rank, world)
    if solve is None:
        if not torch.cuda.is_available():
            raise BaxError("no CUDA device: this package has no CPU fallback")
        solve = ba_main
        kw.setdefault("verbose", False)
    results = {}
    errors = []
    native_ok = (solve is ba_main and os.environ.get("BAX_BATCH_HOST", "0") != "1" and
                 not any(callable(problems[k]) for k in mine) and
                 set(kw) <= {"verbose", "ftol", "xtol", "gtol", "max_nfev", "return_info"} and not kw.get("verbose"))
    if native_ok and mine:
        opts = {k: v for k, v in kw.items() if k != "verbose"}
        results = _ba_batch_native(mine, problems, lambda1, lambda2, n_streams, **opts)
        mine_done = True
    else:
        mine_done = False
    q = queue.SimpleQueue()
    for k in ([] if mine_done else mine):
        q.put(k)

    def worker(i=0):
        cuda = torch.cuda.is_available()
        ctx = torch.cuda.stream(_worker_stream(torch.cuda.current_device(), i)) if cuda else None
        if ctx is not None:
            ctx.__enter__()
        extra = dict(kw)
        reuse = {}
        if solve is ba_main:
            extra["_reuse"] = reuse             # one libbax handle per worker, re-bound to each problem in turn
        try:
            while not errors:
                try:
                    k = q.get_nowait()
                except queue.Empty:
                    return
                p = problems[k]() if callable(problems[k]) else problems[k]
                results[k] = solve(*p, lambda1, lambda2, **extra)
        except BaseException as e:          # surfaced by the caller, with the failing problem's index
            errors.append((k, e))
        finally:
            if reuse.get("h") is not None:
                lib().bax_destroy(reuse.pop("h"))
            if ctx is not None:
                ctx.__exit__(None, None, None)

    n_thr = max(1, min(int(n_streams), len(mine)))
    if mine_done:
        pass
    elif n_thr == 1:
        worker()
    else:
        dev = torch.cuda.current_device() if torch.cuda.is_available() else None

        def entry(i):
            if dev is not None:
                torch.cuda.set_device(dev)
            worker(i)
        threads = [threading.Thread(target=entry, args=(i,), name=f"ba_batch-{i}") for i in range(n_thr)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
    if errors:
        k, e = errors[0]
        raise type(e)(f"ba_batch: problem {k}: {e}") from e
    if group is not None and gather:
        import torch.distributed as dist
        parts = [None] * world
        dist.all_gather_object(parts, results, group=group)
        results = {}
        for part in parts:
            results.update(part)
        return [results[k] for k in range(n)]
    return [results[k] for k in mine]
