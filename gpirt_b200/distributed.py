"""Item-sharded sweeps over several B200s: one process per GPU, torch.distributed (NCCL over NVLink/NVSwitch)
for the exchanges, every kernel and buffer owned by the rank's gpirt_sampler handle.

Given theta, the items are conditionally independent — the reference's `omp parallel for` over j
(src/draw-f.cpp:68, src/draw-fstar.cpp:44, src/draw-beta.cpp:64, src/draw_threshold.cpp:210) — so rank r owns the
contiguous item block [r*m_chunk, (r+1)*m_chunk). The theta step is independent over respondents but needs every
item (src/draw-theta.cpp:123). Default ("items"): it stays sharded BY ITEMS too — every rank scores all respondents
against its own items and the ranks exchange partial log-likelihoods, 16 proposals of every unfinished respondent per
round (include/gpirt_b200.h, gpirt_theta_items_*):

    phase_items   draw_f, draw_fstar on the local items                     (no communication)
    theta rounds  { partials over the local items ; all_gather n x 17 doubles per rank ; decide (same on every rank) }
    phase_close   f <- f* gather, draw_beta, K + Cholesky (replicated), draw_threshold, per-item ll
    all_gather    beta, thresholds, per-item ll packed into one buffer

GPIRT_THETA_SHARD=respondents selects the first scheme of round 1: the theta step sharded by respondents, fed by an all-gather of
the f* shards (8*1001*m bytes per sweep) and a term table over ALL items on every rank.

Random streams are keyed by GLOBAL item / respondent index, so a chain is the same for any world size.
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

    def __init__(self, n, m, Ccat, *, rank, world, device, group=None, theta_shard=None, **kw):
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
        self.theta_shard = os.environ.get("GPIRT_THETA_SHARD", "items") if theta_shard is None else theta_shard
        if self.theta_shard not in ("items", "respondents"):
            raise ValueError(f"theta_shard {self.theta_shard!r}: 'items' or 'respondents'")
        if world == 1:
            self.theta_shard = "respondents"  # one rank: the ordinary kernels
        for name in (("theta_part", "items_packed") if world > 1 else ()):
            ptr, chunk_bytes = self.S.exchange_buffer(name)
            self._bufs[name] = torch.as_tensor(_DevBuf(ptr, chunk_bytes * world), device=f"cuda:{device}")
        self.theta_rounds = 0  # exchange rounds of the last item-sharded theta step
        # config-4 sizes: the theta-only chain (Cholesky of K, L^-1 k*, K^-1 k*) is distributed over the ranks through NVLink peer memory instead of being
        # recomputed on every one of them (include/gpirt_b200.h gpirt_chain_*): every rank maps its peers' arenas
        self.dist_chain = world > 1 and n >= int(os.environ.get("GPIRT_DIST_CHAIN_MIN_N", "4096")) and os.environ.get("GPIRT_DIST_CHAIN", "1") != "0"
        if self.dist_chain:
            import torch.distributed as dist

            exports = [None] * world
            dist.all_gather_object(exports, self.S.chain_export(), group=group)
            for r, ex in enumerate(exports):
                if r != rank:
                    self.S.chain_attach(r, ex)
            dist.barrier(group=group)
        self.p2p = world > 1 and self.theta_shard == "respondents" and os.environ.get("GPIRT_FSTAR_EXCHANGE", "nccl") == "p2p"
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

    def draw_theta(self, z=None, u=None):
        """The theta step alone (parity tests): item-sharded rounds, or the respondent-sharded kernel + theta exchange."""
        if self.theta_shard == "items":
            self.S.theta_items_begin(z=z, u=u)
            self.theta_rounds = 0
            while True:
                self.S.theta_items_partials()
                self._gather("theta_part")
                self.theta_rounds += 1
                if self.S.theta_items_decide() == 0:
                    break
        else:
            if self.p2p:
                self.S.peer_wait()
            else:
                self._gather("fstar")
            self.S.phase_theta()
            self._gather("theta")

    def sweep(self, k=1):
        for _ in range(k):
            self.S.phase_items()
            self.draw_theta()
            self.S.phase_close()
            if self.theta_shard == "items":
                self.S.items_pack()
                self._gather("items_packed")
                self.S.items_pack(unpack=True)
            else:
                self._gather("beta")
                self._gather("thr")
                self._gather("item_ll")

    def total_ll(self):
        return self.S.total_ll()

    def close(self):
        if self.dist_chain:  # no rank frees its arena while a peer still maps it
            import torch.distributed as dist

            self.S.sync()
            self.S.chain_detach()
            dist.barrier(group=self.group)
            self.dist_chain = False
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


class LocalShards:
    """The same item-sharded chain driven from ONE process: `len(devices)` shards, shard r on devices[r] (the same device may appear
    several times), the exchanges done as device-to-device copies of chunk r from shard r into every other shard's buffer (peer copies
    between different GPUs). What the torch.distributed driver does with NCCL, without a process group: for a host application that is one
    process, and for testing the sharded logic on a single GPU. Same calls and same bits as ShardedSampler on `world` ranks."""

    def __init__(self, n, m, Ccat, *, devices, theta_shard="items", **kw):
        import torch

        self.world = len(devices)
        if self.world < 2:
            raise ValueError("LocalShards needs at least two shards; use Sampler for one")
        self.devices, self.theta_shard = list(devices), theta_shard
        self.S, self._bufs = [], []
        for r, dev in enumerate(self.devices):
            S = Sampler(n, m, Ccat, rank=r, world=self.world, device=dev, **kw)
            S.set_stream(torch.cuda.current_stream(dev).cuda_stream)
            bufs = {}
            for name in ("fstar", "beta", "thr", "theta", "item_ll", "theta_part", "items_packed"):
                ptr, chunk_bytes = S.exchange_buffer(name)
                bufs[name] = torch.as_tensor(_DevBuf(ptr, chunk_bytes * self.world), device=f"cuda:{dev}")
            self.S.append(S)
            self._bufs.append(bufs)
        self.theta_rounds = 0
        # the distributed theta-only chain (n >= 4096) needs every shard on a GPU of its own: its exchange points are spinning wait kernels, which on a shared
        # device can be queued in front of the very kernels they wait for
        distinct = len(set(self.devices)) == self.world or os.environ.get("GPIRT_DIST_CHAIN_SAME_DEVICE") == "1"  # the override: experiments only
        self.dist_chain = distinct and n >= int(os.environ.get("GPIRT_DIST_CHAIN_MIN_N", "4096")) and os.environ.get("GPIRT_DIST_CHAIN", "1") != "0"
        if self.dist_chain:
            exports = [S.chain_export() for S in self.S]
            for r, S in enumerate(self.S):
                for q, ex in enumerate(exports):
                    if q != r:
                        S.chain_attach(q, ex, same_process=True)

    def _gather(self, name):
        chunk = self._bufs[0][name].numel() // self.world
        for r in range(self.world):  # chunk r is owned by shard r
            src = self._bufs[r][name][r * chunk:(r + 1) * chunk]
            for q in range(self.world):
                if q != r:
                    self._bufs[q][name][r * chunk:(r + 1) * chunk].copy_(src, non_blocking=True)

    def _each(self, method, *a, **kw):
        return [getattr(S, method)(*a, **kw) for S in self.S]

    def draw_theta(self, z=None, u=None):
        if self.theta_shard == "items":
            self._each("theta_items_begin", z=z, u=u)
            self.theta_rounds = 0
            while True:
                self._each("theta_items_partials")
                self._gather("theta_part")
                self.theta_rounds += 1
                active = self._each("theta_items_decide")
                assert len(set(active)) == 1, f"the shards disagree on the unfinished respondents: {active}"
                if active[0] == 0:
                    break
        else:
            self._gather("fstar")
            self._each("phase_theta")
            self._gather("theta")

    def sweep(self, k=1):
        for _ in range(k):
            self._each("phase_items")
            self.draw_theta()
            self._each("phase_close")
            if self.theta_shard == "items":
                self._each("items_pack")
                self._gather("items_packed")
                self._each("items_pack", unpack=True)
            else:
                for name in ("beta", "thr", "item_ll"):
                    self._gather(name)

    # every shard takes the same full-size arrays and keeps its part
    def set_data(self, y):
        self._each("set_data", y)

    def set_priors(self, *a, **kw):
        self._each("set_priors", *a, **kw)

    def set_state(self, **kw):
        self._each("set_state", **kw)

    def set_iter(self, it):
        self._each("set_iter", it)

    def init_draw(self, **kw):
        self._each("init_draw", **kw)

    def sync(self):
        self._each("sync")

    def get_state(self, **kw):
        """theta / beta / thresholds are complete on every shard after a sweep (shard 0's are returned); f is assembled from the shards' columns."""
        kw.setdefault("fstar", False)
        st = self.S[0].get_state(**kw)
        if kw.get("f", True):
            import ctypes as C

            from ._lib import check

            for S in self.S[1:]:  # gpirt_get_state writes the local items' columns only
                check(S._L.gpirt_get_state(S._h, None, None, st["f"].ctypes.data_as(C.c_void_p), None, None))
        return st

    def diag(self, block):
        ev, mg = self.S[0].diag(block)
        if block != "theta" or self.theta_shard != "items":  # per-unit diagnostics live on the shard that owns the unit
            chunk = self.S[0].n_chunk if block == "theta" else self.S[0].m_chunk
            for r, S in enumerate(self.S[1:], start=1):
                e, g = S.diag(block)
                ev[r * chunk:(r + 1) * chunk], mg[r * chunk:(r + 1) * chunk] = e[r * chunk:(r + 1) * chunk], g[r * chunk:(r + 1) * chunk]
        return ev, mg

    def total_ll(self):
        return self.S[0].total_ll()

    def close(self):
        if self.dist_chain:
            self._each("sync")
            self._each("chain_detach")
        self._each("close")
