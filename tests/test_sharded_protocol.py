"""CPU, world_size 2 and 4 over gloo: the EXCHANGE PROTOCOL of the sharded resample reproduces
the one-piece oracle bit for bit.  Each rank holds its block of the weights and runs, in numpy, exactly the steps the
CUDA library runs per device (slam_b200/csrc/kernels.cuh k_weights_fused + k_search_keep, capi.cu resample_indices):
  exchange 1: per-1024-block totals of w                      -> every rank: ws
  exchange 2: block totals of the in-block CDF and of w^2, and the in-block MAXIMUM of the in-block CDF (+ the 2-slot
              uniform halo)                                    -> every rank: Tc (scanned totals), nEff, BM
  no third exchange: the final CDF is never formed.  Rounding is monotone, so the running maximum of the final CDF
  inside block b is fl(Tc[b-1] + running max of the in-block CDF) and the block maximum is fl(Tc[b-1] + lmax[b]); the
  search adds the prefix to the in-block values it probes in the owner's memory.
  fallback: the exact sequential merge of core.cpp:800-806, the ranks taking turns in rank order.
The library moves the data through peer memory (xchg.cuh pushes the block arrays and the halo, the search reads the
peers' in-block arrays in place); gloo all-gathers stand in for that here — what is checked is WHICH data every rank
must see at each step and that the result is the one-piece oracle's.  The device kernels themselves are checked on
GPUs by tests/test_gpu_sharded.py."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

BLOCK = 1024
f32 = np.float32


def ladder(x):
    """Hillis-Steele inclusive scan over the last axis (length 32), one fp32 add per step and lane."""
    x = x.astype(f32).copy()
    d = 1
    while d < 32:
        y = x.copy()
        y[..., d:] = (x[..., d:] + x[..., :-d]).astype(f32)
        x = y
        d *= 2
    return x


def block_scan(x):
    """Local inclusive scan of consecutive 1024-blocks (zero padded), returns (scan, block totals)."""
    n = x.size
    nb = (n + BLOCK - 1) // BLOCK
    p = np.zeros(nb * BLOCK, f32)
    p[:n] = x
    g = ladder(p.reshape(nb, 32, 32))
    tot = ladder(g[:, :, 31])
    g[:, 1:, :] = (tot[:, :-1, None] + g[:, 1:, :]).astype(f32)
    s = g.reshape(-1)[:n]
    last = np.minimum(np.arange(1, nb + 1) * BLOCK, n) - 1
    return s, s[last].copy()


def upper_scan(tot):
    """The recursive levels over a totals array: returns (scanned totals, blocked sum of the level below)."""
    s, t2 = block_scan(tot)
    if t2.size > 1:
        s2, _ = upper_scan(t2)
        for b in range(1, t2.size):
            s[b * BLOCK:(b + 1) * BLOCK] = (s2[b - 1] + s[b * BLOCK:(b + 1) * BLOCK]).astype(f32)
    total = f32(s[-2] + tot[-1]) if tot.size > 1 else f32(tot[0])
    return s, total


def allgather(x):
    out = [torch.empty_like(torch.from_numpy(x)) for _ in range(dist.get_world_size())]
    dist.all_gather(out, torch.from_numpy(np.ascontiguousarray(x)))
    return np.concatenate([o.numpy() for o in out])


def worker(rank, world, port, P_local, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    Pg = P_local * world
    rng = np.random.default_rng(99)                      # every rank draws the same global inputs, uses its slice
    w_all = (rng.random(Pg) ** 5).astype(f32)
    u_all = rng.random(Pg)
    sl = slice(rank * P_local, (rank + 1) * P_local)
    w = w_all[sl]
    nbl = P_local // BLOCK
    # A / exchange 1 / B: totals of w -> ws
    _, tot = block_scan(w)
    _, ws = upper_scan(allgather(tot))
    # C: normalise, in-block CDF, its running maximum, totals
    wn = (w / ws).astype(f32)
    sq = (wn * wn).astype(f32)
    cdf_loc, tot_c = block_scan(wn)
    _, tot_s = block_scan(sq)
    mloc_loc = np.maximum.accumulate(cdf_loc.reshape(nbl, BLOCK), axis=1).reshape(-1)
    lmax = mloc_loc.reshape(nbl, BLOCK)[:, -1].copy()
    # exchange 2 / D
    tot_c_g, lmax_g = allgather(tot_c), allgather(lmax)
    Tc, _ = upper_scan(tot_c_g)
    _, s2 = upper_scan(allgather(tot_s))
    neff = f32(1.0) / s2
    nbg = Pg // BLOCK
    prefix = np.concatenate([[f32(0)], Tc[:-1]]).astype(f32)
    bmax = np.where(np.arange(nbg) > 0, (prefix + lmax_g).astype(f32), lmax_g)
    BM = np.maximum.accumulate(bmax)
    # what the search reads in the owners' memory: their in-block arrays (peer reads; an all-gather stands in)
    cdf_loc_g, mloc_loc_g = allgather(cdf_loc), allgather(mloc_loc)
    blk = np.arange(Pg) // BLOCK
    key = np.where(blk > 0, (prefix[blk] + mloc_loc_g).astype(f32), mloc_loc_g)          # with_prefix(mloc)
    M = np.where(blk > 0, np.maximum(BM[np.maximum(blk - 1, 0)], key), key)              # running_max_at
    cdf_final = np.where(blk > 0, (prefix[blk] + cdf_loc_g).astype(f32), cdf_loc_g)
    # the identities the kernel relies on
    assert np.array_equal(M, np.maximum.accumulate(cdf_final))
    assert np.array_equal(bmax, np.maximum.reduceat(cdf_final, np.arange(0, Pg, BLOCK)))
    # F: select for local slots (+2 halo from the previous rank), search, keep
    k = f32(1.0 / f32(Pg))
    di = np.zeros(Pg, f32)
    t = f32(k / f32(2))
    for i in range(Pg):
        di[i] = t
        t = f32(t + k)
    lo = max(0, rank * P_local - 2)
    sel = (di[lo:(rank + 1) * P_local].astype(np.float64) + u_all[lo:(rank + 1) * P_local] * np.float64(k)
           - np.float64(f32(k / f32(2)))).astype(f32)
    G = np.searchsorted(M, sel, side="right")        # first index with M > sel
    halo = rank * P_local - lo
    keep = G[halo:].copy()
    prev = np.concatenate([[-1] * (2 - halo), G])[1:1 + P_local] if halo < 2 else G[1:1 + P_local]
    keep = np.maximum(keep, prev)
    keep = np.where(keep >= Pg, -1, keep).astype(np.int32)
    # fallback chain (capi.cu keep_fallback / k_keep_sequential_part): rank r continues where rank r-1 ended
    sel_own = sel[halo:]
    keep_seq = np.full(P_local, -1, np.int32)
    pos = torch.zeros(1, dtype=torch.int64)
    for r in range(world):
        if r == rank:
            i = int(pos.item())
            for c in range(P_local):
                while i < Pg and not (sel_own[c] < cdf_final[i]):
                    i += 1
                keep_seq[c] = i if i < Pg else -1
            pos[0] = i
        dist.broadcast(pos, r)
    q.put((rank, keep, float(neff), wn, keep_seq))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,P_local", [(2, 2048), (4, 1024)])
def test_protocol_equals_one_piece_oracle(oracle, world, P_local):
    port = 29611 + world
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, world, port, P_local, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    Pg = P_local * world
    rng = np.random.default_rng(99)
    w_all = (rng.random(Pg) ** 5).astype(f32)
    u_all = rng.random(Pg)
    wn_o, keep_o, neff_o = oracle.resample_indices(w_all, u_all, scan_mode=1)
    keep = np.concatenate([g[1] for g in got])
    wn = np.concatenate([g[3] for g in got])
    assert np.array_equal(wn, wn_o)
    assert all(g[2] == float(neff_o) for g in got)
    assert np.array_equal(keep, keep_o)
    assert np.array_equal(np.concatenate([g[4] for g in got]), keep_o)   # the sequential chain gives the same keep[]
