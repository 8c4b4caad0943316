"""Logistic regression trained by SSGD / full-batch GD / Model Averaging / BMUF / EASGD.

Reference drivers: ``optimization/ssgd.py:93-112``, ``machine_learning/logistic_regression.py:76-92``,
``optimization/ma.py:93-113``, ``optimization/bmuf.py:98-121``, ``optimization/easgd.py:95-113``.

Three interchangeable backends produce a :class:`LogisticResult`:

``oracle``  NumPy restatement with exact Spark semantics (``oracle.algorithms.logistic_oracle``).
``torch``   plain PyTorch + ``torch.distributed`` (NCCL on GPUs, gloo on CPU) -- the in-box baseline.
``fused``   the sm_100a kernels of ``csrc/lr.cu``: for the reference's small fp64 problem the whole
            training loop is one persistent kernel; at benchmark scale each step is one launch that
            fuses sampling, forward, backward, grid reduction, the cross-GPU all-reduce over peer
            memory and the optimiser update (:class:`ShardedLogisticEngine`).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from .. import _native
from ..config import LogisticConfig
from ..data.datasets import LogisticData
from ..oracle import spark_semantics as sp
from ..oracle.algorithms import LogisticInit, LogisticResult, logistic_oracle
from ..parallel.dist import Dist, current
from ..parallel.partition import RowSharding
from ..utils.trace import span
from ..utils.philox import bernoulli_threshold, philox_row_mask_np, philox_row_mask_torch


# ------------------------------------------------------------------------------------------------
def sample_masks(cfg: LogisticConfig, sizes: np.ndarray, n_iterations: Optional[int] = None) -> Optional[np.ndarray]:
    """Keep-masks [T, N] (uint8) for every iteration, or None for full-batch GD.
    ``spark``: MT19937 stream of ``rdd.sample(False, f, 42 + t)`` (ssgd.py:97); ``philox``: device sampler."""
    T = cfg.n_iterations if n_iterations is None else n_iterations
    if cfg.algo == "gd":
        return None
    n = int(np.sum(sizes))
    if cfg.sampler == "spark":
        return sp.bernoulli_sample_masks(sizes, cfg.mini_batch_fraction, cfg.sample_seed, T).astype(np.uint8)
    return np.stack([philox_row_mask_np(n, cfg.mini_batch_fraction, cfg.sample_seed, t) for t in range(T)]).astype(np.uint8)


def _resolve_backend(backend: str) -> str:
    if backend != "auto":
        return backend
    return "fused" if _native.available() else "torch"


def run_logistic(data: LogisticData, cfg: LogisticConfig, init: Optional[LogisticInit] = None,
                 backend: str = "auto", dist: Optional[Dist] = None,
                 masks: Optional[np.ndarray] = None, record_w: bool = False,
                 checkpoint: Optional[str] = None, checkpoint_every: int = 0, resume: bool = False) -> LogisticResult:
    """Train on ``data`` exactly as the reference script for ``cfg.algo`` does.

    ``checkpoint`` + ``checkpoint_every=k`` write the optimiser state (w, replicas, BMUF momentum, accuracy history)
    every k iterations; ``resume=True`` continues from that file (utils/checkpoint.py).  The resumed trajectory is
    identical to the uninterrupted one because iteration t's mini-batch depends only on (seed, t)."""
    backend = _resolve_backend(backend)
    if init is None:
        init = LogisticInit.draw(cfg.algo, data.dim, cfg.n_slices, cfg.seed)
    if checkpoint and (checkpoint_every > 0 or resume):
        return _run_segmented(data, cfg, init, backend, dist, masks, checkpoint, checkpoint_every, resume)
    if backend == "oracle":
        return logistic_oracle(data.matrix, data.X_test_aug, data.y_test, cfg, init,
                               masks=None if masks is None else masks.astype(bool))
    dist = dist or current()
    sh = RowSharding.spark(data.matrix.shape[0], cfg.n_slices, dist.world)
    if masks is None:
        masks = sample_masks(cfg, sh.sizes())
    if backend == "torch":
        return _run_torch(data, cfg, init, dist, sh, masks)
    if backend == "fused":
        return _run_fused_parity(data, cfg, init, dist, sh, masks, record_w)
    raise ValueError(f"unknown backend {backend!r}")


def _run_segmented(data, cfg, init, backend, dist, masks, path, every, resume) -> LogisticResult:
    """Run ``cfg.n_iterations`` iterations in segments of ``every``, saving a checkpoint after each segment."""
    import dataclasses
    from ..utils import checkpoint as ck
    T = cfg.n_iterations
    if backend != "oracle":
        dist = dist or current()
    is_main = dist is None or dist.is_main
    if masks is None:
        sizes = RowSharding.spark(data.matrix.shape[0], cfg.n_slices, 1 if dist is None else dist.world).sizes()
        masks = sample_masks(cfg, sizes)
    t0, accs, bsz = 0, [], []
    meta = dict(kind="logistic", algo=cfg.algo, dim=int(data.dim), n_slices=int(cfg.n_slices), sampler=cfg.sampler,
                sample_seed=int(cfg.sample_seed))
    if resume:
        import os
        if os.path.exists(path):
            st, m = ck.load_checkpoint(path)
            ck.check_meta(m, **meta)
            t0 = int(m["iteration"])
            init = LogisticInit(w=st["w"], ws=st.get("ws"), delta_w=st.get("delta_w"))
            accs, bsz = st["accs"].tolist(), st["batch_sizes"].tolist()
        elif is_main:
            print(f"resume: no checkpoint at {path}, starting from iteration 0")
    every = every if every > 0 else max(T - t0, 1)
    res = LogisticResult(w=init.w, accs=accs, batch_sizes=bsz, ws=init.ws, delta_w=init.delta_w)
    while t0 < T:
        t1 = min(T, t0 + every)
        seg_cfg = dataclasses.replace(cfg, n_iterations=t1 - t0)
        seg = run_logistic(data, seg_cfg, init, backend, dist, None if masks is None else masks[t0:t1])
        accs, bsz = accs + list(seg.accs), bsz + list(seg.batch_sizes)
        init = LogisticInit(w=seg.w, ws=seg.ws if seg.ws is not None else init.ws, delta_w=seg.delta_w)
        res = LogisticResult(w=seg.w, accs=accs, batch_sizes=bsz, ws=seg.ws, delta_w=seg.delta_w)
        t0 = t1
        if is_main:
            ck.save_checkpoint(path, dict(w=res.w, ws=init.ws, delta_w=init.delta_w, accs=np.asarray(accs, dtype=np.float64),
                                          batch_sizes=np.asarray(bsz, dtype=np.int64)), dict(meta, iteration=t0))
    return res


# ------------------------------------------------------------------------------------------------ torch baseline
def _sigmoid(z: torch.Tensor, eps: float) -> torch.Tensor:
    return 1.0 / (torch.exp(-z) + 1.0 + eps)


def _run_torch(data, cfg, init, dist, sh, masks) -> LogisticResult:
    dev, f64 = dist.device, torch.float64
    parts = sh.local_partitions(dist.rank)
    r0, r1 = sh.local_row_range(dist.rank)
    X = torch.as_tensor(data.X_aug[r0:r1], dtype=f64, device=dev)
    y = torch.as_tensor(data.y[r0:r1], dtype=f64, device=dev)
    loc = sh.local_offsets(dist.rank)
    Xt = torch.as_tensor(data.X_test_aug, dtype=f64, device=dev)
    yt = torch.as_tensor(data.y_test, dtype=torch.int64, device=dev)
    m = None if masks is None else torch.as_tensor(masks[:, r0:r1].astype(np.float64), device=dev)
    D, P, eps = data.dim, cfg.n_slices, cfg.sigmoid_eps
    w = torch.as_tensor(init.w, dtype=f64, device=dev).clone()
    ws = None if init.ws is None else torch.as_tensor(init.ws, dtype=f64, device=dev).clone()
    delta = None if init.delta_w is None else torch.as_tensor(init.delta_w, dtype=f64, device=dev).clone()
    accs, bsz = [], []
    local_algo = cfg.algo in ("ma", "bmuf", "easgd")

    def gather_sum(contrib: torch.Tensor) -> torch.Tensor:
        """contrib [P, D+1] with only local rows filled; partition-ordered sum across ranks (A.3)."""
        dist.all_reduce_sum(contrib)            # rows are disjoint across ranks: this is a gather
        acc = contrib[0].clone()
        for p in range(1, P):
            acc = acc + contrib[p]
        return acc

    for t in range(cfg.n_iterations):
        contrib = torch.zeros(P, D + 1, dtype=f64, device=dev)
        if local_algo and cfg.algo != "easgd":
            ws = w.repeat(P, 1)
        for li, p in enumerate(parts):
            Xp, yp = X[loc[li]:loc[li + 1]], y[loc[li]:loc[li + 1]]
            keep = torch.ones_like(yp) if m is None else m[t, loc[li]:loc[li + 1]]
            cnt = keep.sum()
            if not local_algo:
                r = (_sigmoid(Xp @ w, eps) - yp) * keep
                contrib[p, :D] = r @ Xp
            else:
                wp = ws[p]
                n_inner = 1 if cfg.algo == "easgd" else cfg.n_local_iterations
                if cnt.item() > 0:
                    for _ in range(n_inner):
                        r = (_sigmoid(Xp @ wp, eps) - yp) * keep
                        g_mean = (r @ Xp) / cnt
                        nw = wp - cfg.eta * g_mean
                        if cfg.algo == "easgd":
                            nw = nw - cfg.alpha * (wp - w)
                        wp = nw
                ws[p] = wp
                contrib[p, :D] = wp
            contrib[p, D] = cnt
        S = gather_sum(contrib)
        cnt_tot = float(S[D].item())
        S = S[:D]
        if cfg.algo == "ssgd":
            if cnt_tot > 0:
                reg = _reg_gradient_t(w, cfg.reg_type, cfg.reg_alpha)
                w = w - cfg.eta * (S / cnt_tot + cfg.lam * reg)
        elif cfg.algo == "gd":
            w = w - cfg.eta * S
        elif cfg.algo == "ma":
            w = S / P
        elif cfg.algo == "bmuf":
            delta = cfg.mu * delta + cfg.zeta * (S / P - w)
            w = w + delta
        else:
            w = (1 - cfg.beta) * w + cfg.beta * S / P
        pred = (_sigmoid(Xt @ w, eps) >= 0.5).to(torch.int64)
        accs.append(float((pred == yt).double().mean().item()))
        bsz.append(int(cnt_tot))
    ws_out = None
    if ws is not None:
        full = torch.zeros(P, D, dtype=f64, device=dev)
        for p in parts:
            full[p] = ws[p]
        dist.all_reduce_sum(full)
        ws_out = full.cpu().numpy()
    return LogisticResult(w=w.cpu().numpy(), accs=accs, batch_sizes=bsz, ws=ws_out,
                          delta_w=None if delta is None else delta.cpu().numpy())


def _reg_gradient_t(w, reg_type, alpha):
    if reg_type == "none":
        return torch.zeros_like(w)
    if reg_type == "l2":
        return w
    if reg_type == "l1":
        return torch.sign(w)
    return alpha * torch.sign(w) + (1 - alpha) * w


# ------------------------------------------------------------------------------------------------ fused parity kernel
def _run_fused_parity(data, cfg, init, dist, sh, masks, record_w) -> LogisticResult:
    from ..parallel.symm import default_heap
    lib = _native.load()
    dev, f64 = dist.device, torch.float64
    heap = default_heap()
    parts = sh.local_partitions(dist.rank)
    if not parts:
        raise ValueError("fused parity path needs at least one partition per rank (n_slices >= world)")
    r0, r1 = sh.local_row_range(dist.rank)
    D, P, T = data.dim, cfg.n_slices, cfg.n_iterations
    X = torch.as_tensor(np.ascontiguousarray(data.X_aug[r0:r1]), dtype=f64, device=dev)
    y = torch.as_tensor(np.ascontiguousarray(data.y[r0:r1]), dtype=f64, device=dev)
    off = torch.as_tensor(sh.local_offsets(dist.rank), dtype=torch.int64, device=dev)
    Xt = torch.as_tensor(data.X_test_aug, dtype=f64, device=dev).contiguous()
    yt = torch.as_tensor(data.y_test, dtype=torch.int32, device=dev).contiguous()
    m = None if masks is None else torch.as_tensor(np.ascontiguousarray(masks[:, r0:r1]), dtype=torch.uint8, device=dev)
    w = torch.as_tensor(init.w, dtype=f64, device=dev).clone()
    ws = None
    if cfg.algo in ("ma", "bmuf", "easgd"):
        src = init.ws if init.ws is not None else np.tile(init.w, (P, 1))
        ws = torch.as_tensor(np.ascontiguousarray(src[parts[0]:parts[-1] + 1]), dtype=f64, device=dev).clone()
    delta = None
    if cfg.algo == "bmuf":
        delta = torch.as_tensor(init.delta_w if init.delta_w is not None else np.zeros(D), dtype=f64, device=dev).clone()
    accs = torch.zeros(max(T, 1), dtype=f64, device=dev)
    bsz = torch.zeros(max(T, 1), dtype=torch.int32, device=dev)
    w_hist = torch.zeros(T, D, dtype=f64, device=dev) if record_w else None
    reg = heap.region(f"lr_parity_P{P}_D{D}", int(lib.dapb_lr_parity_payload_bytes(P, D)), P)

    a = _native.DapbLrParityArgs()
    a.X, a.y, a.part_off, a.masks = X.data_ptr(), y.data_ptr(), off.data_ptr(), _native.ptr(m)
    a.X_test, a.y_test = Xt.data_ptr(), yt.data_ptr()
    a.n_local, a.n_test, a.D = r1 - r0, Xt.shape[0], D
    a.P_local, a.P_global, a.part_base, a.T = len(parts), P, parts[0], T
    a.algo, a.n_local_iters = _native.ALGO_ID[cfg.algo], cfg.n_local_iterations
    a.reg_type = _native.REG_ID[cfg.reg_type]
    a.max_part_rows = int(np.max(sh.sizes()))
    a.eta, a.lam, a.reg_alpha, a.mu, a.zeta = cfg.eta, cfg.lam, cfg.reg_alpha, cfg.mu, cfg.zeta
    a.alpha, a.beta, a.eps = cfg.alpha, cfg.beta, cfg.sigmoid_eps
    a.w, a.ws, a.delta_w = w.data_ptr(), _native.ptr(ws), _native.ptr(delta)
    a.accs, a.batch_sizes, a.w_hist = accs.data_ptr(), bsz.data_ptr(), _native.ptr(w_hist)
    a.payload_off, a.flag_base, a.epoch_base = reg.payload_off, reg.flag_base, reg.epoch
    reg.epoch += T
    _native.check(lib.dapb_lr_parity_run(C.byref(a), heap.comm_ref(), _native.stream_ptr()), "lr_parity_run")
    torch.cuda.synchronize()
    # accuracy / batch sizes live on the rank owning partition 0
    if dist.world > 1:
        dist.broadcast(accs, 0)
        dist.broadcast(bsz, 0)
        if w_hist is not None:
            dist.broadcast(w_hist, 0)
    ws_out = None
    if ws is not None:
        full = torch.zeros(P, D, dtype=f64, device=dev)
        full[parts[0]:parts[-1] + 1] = ws
        dist.all_reduce_sum(full)
        ws_out = full.cpu().numpy()
    res = LogisticResult(w=w.cpu().numpy(), accs=accs[:T].cpu().tolist(), batch_sizes=bsz[:T].cpu().tolist(),
                         ws=ws_out, delta_w=None if delta is None else delta.cpu().numpy())
    if w_hist is not None:
        res.w_hist = w_hist.cpu().numpy()
    return res


# ================================================================================================
# Benchmark-scale engine: resident shard, one fused launch per optimiser step
# ================================================================================================
_DTYPE_ID = {torch.float32: 0, torch.bfloat16: 1}


class ShardedLogisticEngine:
    """Per-rank engine over a resident ``[n_rows, D]`` shard (bias column included).

    One logical partition / replica per rank (``n_slices == world``: BASELINE.json config 2,
    "8 workers on 8xB200").  ``backend='fused'`` launches ``dapb_lr_step`` once per local step;
    ``backend='nccl'`` is the straightforward ``torch.matmul`` + ``dist.all_reduce`` baseline fed the
    same Philox masks.
    """

    def __init__(self, X: torch.Tensor, y: torch.Tensor, cfg: LogisticConfig, row_base: int = 0,
                 dist: Optional[Dist] = None, backend: str = "fused", w0: Optional[torch.Tensor] = None,
                 grid: int = 0, heap=None, use_tc: bool = False, collect_stats: bool = False):
        self.dist = dist or current()
        self.cfg, self.backend = cfg, backend
        assert X.dim() == 2 and X.is_contiguous() and X.dtype in _DTYPE_ID
        self.X, self.y = X, y.float().contiguous()
        self.n_rows, self.D = X.shape
        self.row_base = int(row_base)
        dev = X.device
        g = torch.Generator(device="cpu").manual_seed(cfg.seed if cfg.seed is not None else 0)
        if w0 is None:
            w0 = (2 * torch.rand(self.D, generator=g) - 1) * 0.01
        self.w = w0.to(device=dev, dtype=torch.float32).clone().contiguous()
        self.w_local = self.w.clone()
        self.delta_w = torch.zeros_like(self.w)
        self.t = 0
        self.P = self.dist.world
        self.threshold24 = (1 << 24) if cfg.algo == "gd" else bernoulli_threshold(cfg.mini_batch_fraction)
        self.last_count = torch.zeros(1, dtype=torch.int32, device=dev)
        if backend == "fused":
            from ..parallel.symm import default_heap
            self.lib = _native.load()
            self.heap = heap or default_heap()
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            # tensor-core variant (csrc/lr.cu::lr_tc_kernel): bf16 rows with D == 1024, one CTA per SM
            self.use_tc = bool(use_tc) and X.dtype == torch.bfloat16 and self.D == 1024
            self.grid = grid or (sms if self.use_tc else 2 * sms)
            self.partials = torch.empty(self.grid, self.D, dtype=torch.float32, device=dev)
            self.counts = torch.zeros(self.grid, dtype=torch.int32, device=dev)
            self.ticket = torch.zeros(1, dtype=torch.int32, device=dev)
            self.g = torch.zeros(self.D, dtype=torch.float32, device=dev)
            nbytes = int(self.lib.dapb_lr_step_payload_bytes(_native.MAX_RANKS, self.D))
            self.region = self.heap.region(f"lr_step_D{self.D}", nbytes, _native.LR_STEP_FLAGS)
            self.samples_seen = torch.zeros(1, dtype=torch.int64, device=dev)      # running sum of batch sizes (in-kernel)
            self.stats = torch.zeros(4, dtype=torch.int64, device=dev) if collect_stats else None
            self._eval_scratch = torch.zeros(2, dtype=torch.int32, device=dev)
            self._acc = torch.zeros(1, dtype=torch.float32, device=dev)
            self._args = self._base_args()

    # ------------------------------------------------------------------ fused
    def _base_args(self) -> _native.DapbLrStepArgs:
        cfg, a = self.cfg, _native.DapbLrStepArgs()
        a.X, a.y, a.n_rows, a.row_base = self.X.data_ptr(), self.y.data_ptr(), self.n_rows, self.row_base
        a.D, a.dtype, a.mask = self.D, _DTYPE_ID[self.X.dtype], 0
        a.seed, a.threshold24, a.eps = cfg.sample_seed, self.threshold24, cfg.sigmoid_eps
        a.partials, a.counts, a.ticket, a.grid = self.partials.data_ptr(), self.counts.data_ptr(), self.ticket.data_ptr(), self.grid
        a.algo, a.reg_type, a.use_sum = _native.ALGO_ID[cfg.algo], _native.REG_ID[cfg.reg_type], int(cfg.algo == "gd")
        a.eta, a.lam, a.reg_alpha, a.mu, a.zeta = cfg.eta, cfg.lam, cfg.reg_alpha, cfg.mu, cfg.zeta
        a.alpha = cfg.eta * cfg.rho
        a.beta = self.P * cfg.eta * cfg.rho
        a.n_replicas_total = self.P
        a.g_out, a.cnt_out = self.g.data_ptr(), self.last_count.data_ptr()
        a.w, a.w_local, a.delta_w = self.w.data_ptr(), self.w_local.data_ptr(), self.delta_w.data_ptr()
        a.payload_off, a.flag_base = self.region.payload_off, self.region.flag_base
        a.cnt_accum = self.samples_seen.data_ptr()
        a.stats = self.stats.data_ptr() if self.stats is not None else 0
        return a

    def _launch(self, fin: int, w_fwd: torch.Tensor, iteration: int, reset: int = 0, stream=None):
        a = self._args
        a.fin, a.w_fwd, a.iteration, a.reset_replica = fin, w_fwd.data_ptr(), iteration, reset
        if fin in (_native.FIN_SSGD, _native.FIN_LOCAL_SYNC):
            a.epoch = self.region.next_epoch()
        fn = self.lib.dapb_lr_step_tc if self.use_tc else self.lib.dapb_lr_step
        _native.check(fn(C.byref(a), self.heap.comm_ref(), _native.stream_ptr(stream)), "lr_step")

    def state_dict(self) -> dict:
        """Optimiser state: global ``w`` and BMUF momentum are replicated; ``w_local`` (the EASGD replica) is
        rank-local and belongs in ``checkpoint.rank_path(path, rank)``.  The mini-batch of iteration t is a pure
        function of (seed, t, row), so ``t`` is all the sampler state there is."""
        return {"w": self.w.detach().cpu().numpy(), "w_local": self.w_local.detach().cpu().numpy(),
                "delta_w": self.delta_w.detach().cpu().numpy(), "t": np.int64(self.t)}

    def load_state_dict(self, st: dict):
        for name in ("w", "w_local", "delta_w"):
            getattr(self, name).copy_(torch.as_tensor(np.asarray(st[name]), dtype=torch.float32))
        self.t = int(st["t"])

    def step(self, stream=None):
        """One outer iteration / communication round of ``cfg.algo`` (ref loop bodies cited above)."""
        with span("logistic.step"):
            self._step(stream)
        self.t += 1

    def _step(self, stream=None):
        cfg, t = self.cfg, self.t
        if self.backend == "nccl":
            self._step_nccl()
        elif cfg.algo in ("ssgd", "gd"):
            self._launch(_native.FIN_SSGD, self.w, t, stream=stream)
        else:
            n_inner = 1 if cfg.algo == "easgd" else cfg.n_local_iterations
            for li in range(n_inner):                     # same mini-batch every local step (ma.py:99)
                last = li == n_inner - 1
                self._launch(_native.FIN_LOCAL_SYNC if last else _native.FIN_LOCAL, self.w_local, t,
                             reset=int(cfg.algo != "easgd"), stream=stream)

    def evaluate(self, X_test: torch.Tensor, y_test: torch.Tensor) -> torch.Tensor:
        """Accuracy of the current global ``w`` on a (replicated) test set; stays on device."""
        if self.backend == "fused":
            _native.check(self.lib.dapb_lr_eval(X_test.data_ptr(), y_test.data_ptr(), X_test.shape[0], self.D,
                                                _DTYPE_ID[X_test.dtype], self.w.data_ptr(), float(self.cfg.sigmoid_eps),
                                                self._acc.data_ptr(), self._eval_scratch.data_ptr(),
                                                _native.stream_ptr()), "lr_eval")
            return self._acc
        z = X_test.float() @ self.w
        pred = (1.0 / (torch.exp(-z) + 1.0 + self.cfg.sigmoid_eps) >= 0.5).float()
        return (pred == y_test.float()).float().mean().reshape(1)

    # ------------------------------------------------------------------ NCCL baseline
    def _grad_nccl(self, w: torch.Tensor, keep: torch.Tensor):
        idx = keep.nonzero(as_tuple=False).squeeze(1)
        Xs = self.X.index_select(0, idx)
        ys = self.y.index_select(0, idx)
        z = (Xs @ w.to(Xs.dtype)).float()
        r = 1.0 / (torch.exp(-z) + 1.0 + self.cfg.sigmoid_eps) - ys
        g = (r.to(Xs.dtype) @ Xs).float()
        return g, idx.numel()

    def _keep_mask_nccl(self) -> torch.Tensor:
        """The step's Bernoulli mask for the baseline.  On a GPU it comes from the library's one-pass Philox mask
        kernel (33 M rows in tens of microseconds), so the baseline's step time is its gather + two GEMVs +
        all-reduce and not an emulated RNG; on CPU (gloo tests) the torch int64 emulation is used."""
        cfg = self.cfg
        if cfg.algo == "gd":
            return torch.ones(self.n_rows, dtype=torch.bool, device=self.X.device)
        if self.X.is_cuda and _native.available():
            if getattr(self, "_mask_buf", None) is None or self._mask_buf.numel() < self.n_rows:
                self._mask_buf = torch.empty(self.n_rows, dtype=torch.uint8, device=self.X.device)
            _native.check(_native.load().dapb_philox_mask(self.row_base, self.n_rows, cfg.sample_seed, self.t,
                                                          self.threshold24, self._mask_buf.data_ptr(),
                                                          _native.stream_ptr()), "philox_mask")
            return self._mask_buf[:self.n_rows].view(torch.bool)
        return philox_row_mask_torch(self.n_rows, cfg.mini_batch_fraction, cfg.sample_seed, self.t,
                                     self.row_base, device=self.X.device)

    def _step_nccl(self):
        cfg, d = self.cfg, self.dist
        keep = self._keep_mask_nccl()
        if cfg.algo in ("ssgd", "gd"):
            g, cnt = self._grad_nccl(self.w, keep)
            buf = torch.cat([g, torch.full((1,), float(cnt), device=g.device)])
            d.all_reduce_sum(buf)
            g, cnt_tot = buf[:-1], buf[-1]                         # stays on the device: no host round trip
            if cfg.algo == "gd":
                self.w -= cfg.eta * g
            else:
                upd = cfg.eta * (g / cnt_tot.clamp(min=1.0) + cfg.lam * _reg_gradient_t(self.w, cfg.reg_type, cfg.reg_alpha))
                self.w -= torch.where(cnt_tot > 0, upd, torch.zeros_like(upd))
            self.last_count.copy_(cnt_tot.to(torch.int32).reshape(1))
            return
        alpha, beta = cfg.eta * cfg.rho, self.P * cfg.eta * cfg.rho
        if cfg.algo != "easgd":
            self.w_local.copy_(self.w)
        n_inner = 1 if cfg.algo == "easgd" else cfg.n_local_iterations
        for _ in range(n_inner):
            g, cnt = self._grad_nccl(self.w_local, keep)
            self.last_count.fill_(cnt)
            if cnt > 0:
                nw = self.w_local - cfg.eta * g / cnt
                if cfg.algo == "easgd":
                    nw = nw - alpha * (self.w_local - self.w)
                self.w_local.copy_(nw)
        S = self.w_local.clone()
        d.all_reduce_sum(S)
        if cfg.algo == "ma":
            self.w.copy_(S / self.P)
        elif cfg.algo == "bmuf":
            self.delta_w.copy_(cfg.mu * self.delta_w + cfg.zeta * (S / self.P - self.w))
            self.w += self.delta_w
        else:
            self.w.copy_((1 - beta) * self.w + beta * S / self.P)


class StreamingLogisticEngine:
    """End-to-end streaming trainer: every step's mini-batch arrives from *pinned host memory*.

    ``fit_stream(batches)`` uploads batch ``i + 1`` on a copy stream while batch ``i`` runs through the
    fused step kernel (all rows of the batch are used; the data loader did the sampling), evaluates the
    updated model on a resident test set -- the reference computes test accuracy every iteration
    (optimization/ssgd.py:107-112) -- and yields ``(accuracy, global batch size)`` read back to the host.
    """

    def __init__(self, dim: int, dtype: torch.dtype, max_batch_rows: int, cfg: LogisticConfig,
                 X_test: torch.Tensor, y_test: torch.Tensor, dist: Optional[Dist] = None, backend: str = "fused"):
        self.dist = dist or current()
        dev = self.dist.device
        self.bufs = [(torch.empty(max_batch_rows, dim, dtype=dtype, device=dev),
                      torch.empty(max_batch_rows, dtype=torch.float32, device=dev)) for _ in range(2)]
        gd_like = LogisticConfig(**{**cfg.__dict__})
        gd_like.mini_batch_fraction = 1.0
        self.engines = [ShardedLogisticEngine(Xb, yb, gd_like, dist=self.dist, backend=backend) for Xb, yb in self.bufs]
        self.engines[0].threshold24 = self.engines[1].threshold24 = 1 << 24      # use every uploaded row
        for e in self.engines:
            if backend == "fused":
                e._args.threshold24 = 1 << 24
        # both engines share one set of weights
        self.engines[1].w, self.engines[1].w_local, self.engines[1].delta_w = (self.engines[0].w, self.engines[0].w_local,
                                                                                self.engines[0].delta_w)
        if backend == "fused":
            a = self.engines[1]._args
            a.w, a.w_local, a.delta_w = self.engines[0].w.data_ptr(), self.engines[0].w_local.data_ptr(), self.engines[0].delta_w.data_ptr()
        self.X_test, self.y_test = X_test, y_test.float()
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.result_host = torch.zeros(2, dtype=torch.float32).pin_memory()
        self.result_dev = torch.zeros(2, dtype=torch.float32, device=dev)
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    @property
    def w(self):
        return self.engines[0].w

    def _upload(self, slot: int, Xh: torch.Tensor, yh: torch.Tensor):
        Xd, yd = self.bufs[slot]
        n = Xh.shape[0]
        with torch.cuda.stream(self.copy_stream):
            Xd[:n].copy_(Xh, non_blocking=True)
            yd[:n].copy_(yh, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        self.h2d_bytes += Xh.numel() * Xh.element_size() + yh.numel() * yh.element_size()
        return n, ev

    def fit_stream(self, batches):
        it = iter(batches)
        try:
            nxt = next(it)
        except StopIteration:
            return
        slot = 0
        pending = self._upload(slot, *nxt)
        main = torch.cuda.current_stream()
        done_compute = [None, None]
        while pending is not None:
            n, ev = pending
            cur = slot
            try:
                nxt = next(it)
                slot ^= 1
                if done_compute[slot] is not None:
                    self.copy_stream.wait_event(done_compute[slot])     # buffer free again?
                pending = self._upload(slot, *nxt)
            except StopIteration:
                pending = None
            main.wait_event(ev)
            eng = self.engines[cur]
            eng.n_rows = n
            if eng.backend == "fused":
                eng._args.n_rows = n
            eng.step()
            acc = eng.evaluate(self.X_test, self.y_test)
            self.result_dev[0:1].copy_(acc)
            self.result_dev[1:2].copy_(eng.last_count.float())
            done = torch.cuda.Event()
            done.record(main)
            done_compute[cur] = done
            self.result_host.copy_(self.result_dev, non_blocking=True)
            main.synchronize()                                           # the step's result is on the host
            self.d2h_bytes += 8
            yield float(self.result_host[0]), int(self.result_host[1])
