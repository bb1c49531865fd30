"""World-size-2 gloo tests (CPU) of the multi-process drivers in pcoll_b200.sharding: the partitioning
and the collectives' plumbing.  The compute is a deterministic stand-in context defined here (test
infrastructure; the product has no CPU path): what is asserted is that the sharded job equals the
single-process job and that every rank ends with the same result."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import pcoll_b200
from pcoll_b200 import sharding


class StubAbi:
    MODE_PHILOX = 1

    @staticmethod
    def make_mc_params(plan, n_rollouts=0, rollout_offset=0, **kw):
        return dict(n=n_rollouts, offset=rollout_offset, T=plan["trajectory"].shape[1]), []


class StubCtx:
    """Counter-based stand-in: rollout id i 'hits' at waypoint i % (2 T) when that is < T."""
    device = None

    def mc_prepare(self, p):
        self.T = p["T"]
        self.counters = torch.zeros(4 + self.T, dtype=torch.int64)

    def mc_launch(self, n, offset):
        ids = np.arange(offset, offset + n, dtype=np.int64)
        first = ids % (2 * self.T)
        hit = first < self.T
        self.counters[0] += int(hit.sum())
        self.counters[2] += n * self.T
        self.counters[3] += n * self.T * 63
        self.counters[4:] += torch.from_numpy(np.bincount(first[hit], minlength=self.T))

    def mc_counters_tensor(self, T):
        return self.counters

    def mc_collect(self, T, want_hist=False):
        c = self.counters.numpy()
        return dict(hits=int(c[0]), near_contact=int(c[1]), executed_steps=int(c[2]), pair_tests=int(c[3]),
                    first_hit_hist=c[4:].astype(np.uint64).copy())

    def score_plans(self, model, tr, od, rollouts, plan_offset=0, **kw):
        ids = plan_offset + np.arange(tr.shape[0])
        return dict(P=(ids * 37 % 101) / 101.0)

    # EM stand-in: plain numpy E-step statistics about the current mean, textbook M-step
    def em_prepare(self, x, hefts, means, fcovs, iters, var_floor=1e-10, n_total=0):
        self.x = np.asarray(x, dtype=np.float64)
        self.h, self.m, self.c = (np.array(v, dtype=np.float64) for v in (hefts, means, fcovs))
        self.n_total = n_total
        F, K = self.h.shape
        self.stats = torch.zeros(F * (K * 10 + 1), dtype=torch.float64)

    def em_estep(self):
        F, K = self.h.shape
        st = self.stats.view(F, K * 10 + 1)
        iu = np.triu_indices(3)
        for f in range(F):
            X = self.x[f].T
            ll = np.empty((X.shape[0], K))
            for g in range(K):
                d = X - self.m[f, g]
                A = np.linalg.inv(self.c[f, g])
                ll[:, g] = -0.5 * np.einsum("ni,ij,nj->n", d, A, d) - 0.5 * np.log(np.linalg.det(self.c[f, g])) \
                    - 1.5 * np.log(2 * np.pi) + np.log(self.h[f, g])
            mx = ll.max(1, keepdims=True)
            lse = mx[:, 0] + np.log(np.exp(ll - mx).sum(1))
            r = np.exp(ll - lse[:, None])
            for g in range(K):
                d = X - self.m[f, g]
                dd = np.einsum("n,ni,nj->ij", r[:, g], d, d)
                st[f, g * 10:(g + 1) * 10] += torch.from_numpy(np.concatenate([[r[:, g].sum()], r[:, g] @ d, dd[iu]]))
            st[f, K * 10] += float(lse.sum())

    def em_stats_tensor(self):
        return self.stats

    def em_mstep(self):
        F, K = self.h.shape
        st = self.stats.view(F, K * 10 + 1).numpy()
        iu = np.triu_indices(3)
        for f in range(F):
            for g in range(K):
                s = st[f, g * 10:(g + 1) * 10]
                delta = s[1:4] / s[0]
                C = np.zeros((3, 3)); C[iu] = s[4:] / s[0]; C = C + C.T - np.diag(np.diag(C))
                self.c[f, g] = C - np.outer(delta, delta)
                self.m[f, g] = self.m[f, g] + delta
                self.h[f, g] = s[0] / self.n_total
        self.stats.zero_()

    def em_collect(self):
        return dict(hefts=self.h.copy(), means=self.m.copy(), fcovs=self.c.copy())


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _job(world):
    """Everything one rank computes; world == 1 without a process group is the single-process answer."""
    plan = dict(trajectory=np.zeros((3, 56)))
    out = {}
    mc = sharding.mc_convergence_run(StubCtx(), StubAbi, plan, 100003)
    out["mc"] = (mc["hits"], mc["executed_steps"], mc["pair_tests"], mc["first_hit_hist"].tolist(), mc["P"])
    tr, od = np.zeros((37, 3, 20)), np.zeros((37, 3, 19))
    out["score"] = sharding.score_plans_sharded(StubCtx(), {}, tr, od, 1024).tolist()
    x, h, m, c = pcoll_b200.inputs.synthetic_em_problem(3, 4001, 3)
    rank, w = (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)
    lo, hi = sharding.shard_range(x.shape[1], rank, w)
    em = sharding.em_fit_sharded(StubCtx(), x[None, :, lo:hi], x.shape[1], h[None], m[None], c[None], 4)
    out["em"] = (em["hefts"].tolist(), em["means"].tolist(), em["fcovs"].tolist())
    return out


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q.put((rank, _job(world)))
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions():
    for total in (0, 1, 7, 100003, 4096):
        for world in (1, 2, 3, 8):
            r = [sharding.shard_range(total, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == total
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


@pytest.mark.timeout(180)
def test_two_rank_gloo_job_equals_single_process():
    single = _job(1)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=150) for _ in range(2))
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    for rank in (0, 1):
        assert got[rank]["mc"] == single["mc"]
        assert got[rank]["score"] == single["score"]
        for a, b in zip(got[rank]["em"], single["em"]):
            np.testing.assert_allclose(np.array(a), np.array(b), rtol=1e-9, atol=1e-12)
    assert got[0]["em"] == got[1]["em"]
