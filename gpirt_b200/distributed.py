"""Item-sharded sweeps over several B200s: one process per GPU, torch.distributed (NCCL over NVLink/NVSwitch)
for the exchanges, every kernel and buffer owned by the rank's gpirt_sampler handle.

Given theta, the items are conditionally independent — the reference's `omp parallel for` over j
(src/draw-f.cpp:68, src/draw-fstar.cpp:44, src/draw-beta.cpp:64, src/draw_threshold.cpp:210) — so rank r owns the
contiguous item block [r*m_chunk, (r+1)*m_chunk). The theta step is independent over respondents but needs every
item (src/draw-theta.cpp:123), so it is sharded by respondents and fed by one all-gather of the f* shards:

    phase_items   draw_f, draw_fstar on the local items                     (no communication)
    all_gather    f* chunks (8*1001*m bytes in total), beta, thresholds
    phase_theta   theta ESS for the local respondents over the whole table
    all_gather    theta (8*n bytes)
    phase_close   f <- f* gather, draw_beta, K + Cholesky (replicated), draw_threshold, per-item ll
    all_gather    beta, thresholds, per-item ll  (tiny)

Random streams are keyed by GLOBAL item / respondent index, so a chain is bit-identical for any world size.
"""
import os

import numpy as np

from .sampler import Sampler


def shard_plan(n, m, world):
    """Partition used by the C-ABI (gpirt_create): equal ceil-sized chunks, the last one ragged."""
    m_chunk, n_chunk = (m + world - 1) // world, (n + world - 1) // world
    plan = []
    for r in range(world):
        i0, r0 = r * m_chunk, r * n_chunk
        plan.append(dict(rank=r, item0=i0, m_local=max(0, min(m_chunk, m - i0)), resp0=r0, n_local=max(0, min(n_chunk, n - r0))))
    return dict(m_chunk=m_chunk, n_chunk=n_chunk, ranks=plan)


def all_gather_chunks(buf, rank, world, group=None):
    """In-place all-gather of a flat tensor made of `world` equal chunks, chunk r owned by rank r."""
    import torch.distributed as dist

    chunk = buf.numel() // world
    assert chunk * world == buf.numel()
    if world == 1:
        return
    mine = buf[rank * chunk:(rank + 1) * chunk]
    if buf.is_cuda:
        dist.all_gather_into_tensor(buf, mine, group=group)
    else:  # gloo (CPU tests of the host logic)
        parts = [buf[r * chunk:(r + 1) * chunk] for r in range(world)]
        dist.all_gather(parts, mine.clone(), group=group)


class _DevBuf:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes // 8,), "typestr": "<f8", "data": (ptr, False), "version": 2}


class ShardedSampler:
    """One rank's shard of a chain. Every rank passes the same full-size y / state arrays; the handle keeps its part."""

    def __init__(self, n, m, Ccat, *, rank, world, device, group=None, **kw):
        import torch

        self.rank, self.world, self.group = rank, world, group
        self.S = Sampler(n, m, Ccat, rank=rank, world=world, device=device, **kw)
        self.S.set_stream(torch.cuda.current_stream(device).cuda_stream)
        self._bufs = {}
        for name in ("fstar", "beta", "thr", "theta", "item_ll"):
            ptr, chunk_bytes = self.S.exchange_buffer(name)
            self._bufs[name] = torch.as_tensor(_DevBuf(ptr, chunk_bytes * world), device=f"cuda:{device}")
        # The f* exchange (8 x 1001 x m bytes per sweep, the only large one) has two implementations with the same bits (it moves finished
        # values): the in-place NCCL all-gather (default), and GPIRT_FSTAR_EXCHANGE=p2p — every rank maps its peers' arenas (cudaIpc) and the
        # last kernel of the f* draw stores into all of them over NVLink (gpirt_ipc_*, include/gpirt_b200.h). Measured at 8 x B200 on C3 blocks
        # (profiles/r01_scaling.txt): the push costs 0.41 ms (280 MB per GPU at 680 GB/s, the NVLink egress floor) against ~0.45 ms for NCCL,
        # 6.80 vs 6.86 ms per sweep — but mapping 7 peers costs ~200 ms per handle, which a 20-sweep gpirtMCMC call never earns back, hence opt-in.
        self.p2p = world > 1 and os.environ.get("GPIRT_FSTAR_EXCHANGE", "nccl") == "p2p"
        if self.p2p:
            import torch.distributed as dist

            handles = [None] * world
            dist.all_gather_object(handles, self.S.ipc_export(), group=group)
            for r, hnd in enumerate(handles):
                if r != rank:
                    self.S.ipc_attach(r, hnd)
            dist.barrier(group=group)

    def _gather(self, name):
        all_gather_chunks(self._bufs[name], self.rank, self.world, self.group)

    def sweep(self, k=1):
        for _ in range(k):
            self.S.phase_items()
            if self.p2p:
                self.S.peer_wait()
            else:
                self._gather("fstar")
            self.S.phase_theta()
            self._gather("theta")
            self.S.phase_close()
            self._gather("beta")
            self._gather("thr")
            self._gather("item_ll")

    def total_ll(self):
        return self.S.total_ll()

    def close(self):
        if self.p2p:  # no rank frees its arena while a peer still maps it
            import torch.distributed as dist

            self.S.ipc_detach()
            dist.barrier(group=self.group)
            self.p2p = False
        self.S.close()

    def __getattr__(self, name):  # set_data, set_priors, set_state, get_state, sync, ...
        return getattr(self.S, name)


def gpirtMCMC_sharded(data, sample_iterations, burn_iterations, THIN=1, *, rank, world, device, group=None, beta_prior_sds=None,
                      theta_prior_sds=None, theta_os=1, theta_ls=10, KERNEL="Matern", theta_init=None, thresholds=None, SEED=1, chain=0,
                      beta_init=None, f_init=None):
    """gpirtMCMC() (R/gpirtMCMC.R:98-167, src/gpirtMCMC.cpp:25-406) for one chain sharded over `world` ranks: every rank calls it with
    the same host arrays and gets the same theta / beta / threshold / ll draws back (reference layouts). Each rank uploads only its
    items' columns and its respondents' rows of y; kept draws are buffered on the device and read back in bulk. f / f* draws are
    not kept (each rank holds its own items only). The chain is the one the single-GPU gpirtMCMC(SEED=SEED) produces."""
    from .api import _as_cube, _F

    y = _as_cube(data)
    n, m, H = y.shape
    if H != 1:
        raise NotImplementedError("horizon > 1")
    thr = np.asfortranarray(np.asarray(thresholds, dtype=np.float64).reshape(m, -1))
    Ccat = thr.shape[1] - 1
    bps = _F(np.full((3, m), 3.0) if beta_prior_sds is None else beta_prior_sds)
    tps = _F(np.zeros((2, n)) if theta_prior_sds is None else theta_prior_sds)
    S = ShardedSampler(n, m, Ccat, rank=rank, world=world, device=device, group=group, theta_os=theta_os, theta_ls=theta_ls, KERNEL=KERNEL,
                       seed=SEED, chain=chain)
    try:
        S.set_data(y[:, :, 0])
        S.set_priors(bps, tps)
        S.set_state(theta=theta_init, beta=beta_init, f=f_init, thresholds=thr)
        S.set_iter(0xFFFFFFFF)  # as gpirt_mcmc: the initial draws use a sweep id no sweep reaches
        S.init_draw(draw_beta=beta_init is None, draw_f=f_init is None)
        S.set_iter(0)
        n_samples = int(sample_iterations / THIN)  # src/gpirtMCMC.cpp:48
        out = dict(theta=np.zeros((n_samples, n, 1), order="F"), beta=np.zeros((3, m, 1, n_samples), order="F"),
                   threshold=np.zeros((m, Ccat + 1, 1, n_samples), order="F"), ll=np.zeros(n_samples))
        if n_samples > 0:
            S.store_open(n_samples)
        for it in range(sample_iterations + burn_iterations):
            S.sweep(1)
            if it >= burn_iterations and it % THIN == 0 and (it - burn_iterations) // THIN < n_samples:  # :334-335
                S.store_push()
        if n_samples > 0:
            S.store_flush(out, 0)
    finally:
        S.close()
    return dict(theta=out["theta"], beta=[out["beta"][..., s] for s in range(n_samples)],
                threshold=[out["threshold"][..., s] for s in range(n_samples)], ll=out["ll"], f=None, fstar=None)
