"""ngs_schedule_batch_noalloc — the batch path for hosts that may not allocate device memory: an
extender living inside a process whose device arena belongs to another allocator (a cudaMalloc of
ours could fail or fragment it), or one that must keep cudaMalloc's device-wide synchronisation
off the scheduling path.  The library brings its own fixed scratch for that call; tables, pods
and outputs are the caller's (here: torch tensors).

Checked through the C ABI against the existing handle-based entry point (ngs_schedule_batch_dev)
on the same snapshot and pods, bit for bit, on two calls in a row — the second call finds the
scratch array full of the first call's words — and, so that the comparison does not rest on the
handle path alone, the first call's decisions also against the CPU oracle.  This file sorts last
on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NOW_S = 1_790_000_000
NOW_NS = NOW_S * 10**9 + 5


def _tables_on_device(torch, gcount, free, load_tables):
    """the caller-owned snapshot, encoded as include/ngs_dealer.h documents"""
    n = len(gcount)
    dev = torch.device("cuda", 0)
    t = {
        "free": torch.from_numpy(free.astype(np.int32)).to(dev),
        "gpu_count": torch.from_numpy(gcount.astype(np.uint8)).to(dev),
        "remain_load": torch.zeros((n, 8), dtype=torch.int8, device=dev),
        "plan_tag": torch.zeros(n, dtype=torch.int32, device=dev),
        "plan_score": torch.zeros(n, dtype=torch.int32, device=dev),
    }
    tenths = np.full((2, n, 8), 255, dtype=np.uint8)
    ts = np.zeros((2, n, 8), dtype=np.int32)
    if load_tables is not None:
        for m, (value, stamp, absent) in enumerate(load_tables):
            ok = ~absent[:, None] & (value >= 0) & (value <= 1)
            tenths[m] = np.where(ok, np.ceil(10 * value), 255).astype(np.uint8)
            ts[m] = np.where(ok, stamp, 0).astype(np.int32)
    t["tenths"] = torch.from_numpy(tenths).to(dev)
    t["ts"] = torch.from_numpy(ts).to(dev)
    return t


def _view(ngs, t, n):
    v = ngs._abi.SnapshotView()
    for k in ("free", "gpu_count", "remain_load", "tenths", "ts", "plan_tag", "plan_score"):
        setattr(v, k, t[k].data_ptr())
    v.n_nodes = n
    v.first_node = 0
    return v


@pytest.mark.parametrize("load_on", [False, True])
def test_noalloc_matches_handle_path_twice(built, load_on):
    import torch

    import nano_gpu_scheduler_b200 as ngs
    from nano_gpu_scheduler_b200 import workloads as wl

    n, P, max_c = 6000, 200, 4
    dev = torch.device("cuda", 0)
    gcount, free = wl.make_cluster(n, 0.5, 0x2A)
    load_tables = wl.make_load_tables(n, 0x2A, NOW_S) if load_on else None
    load = ngs.make_load_ctx(wl.DEFAULT_PERIODS, NOW_NS) if load_on else None

    t = _tables_on_device(torch, gcount, free, load_tables)
    view = _view(ngs, t, n)
    with ngs.GpuDealer(n, device=0) as ref:
        ref.snapshot_load(gcount, free)
        if load_on:
            for m, (value, stamp, absent) in enumerate(load_tables):
                ok = np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8)
                ref.load_update_bulk(m, value, stamp, value_ok=ok, ts_ok=ok)
        for call in range(2):  # two calls in a row on the same tables
            pc, nc = wl.make_pods(P, 1, 4, 0, 12, 0x2A + call, max_c=max_c)
            pol = ((np.arange(P) + call) % 3).astype(np.int32)  # binpack / spread / random interleaved
            d_pc, d_nc, d_pol = (torch.from_numpy(a).to(dev) for a in (pc, nc, pol))
            want = torch.empty((2 + max_c) * P, dtype=torch.int32, device=dev)
            got = torch.full(((2 + max_c) * P,), 12345, dtype=torch.int32, device=dev)
            ref.schedule_batch_dev(P, d_pc.data_ptr(), d_nc.data_ptr(), max_c, d_pol.data_ptr(), want.data_ptr(),
                                   want.data_ptr() + 4 * P, want.data_ptr() + 8 * P, load=load, seed=9, seq0=call * P)
            ref.sync()
            torch.cuda.synchronize()
            rc = ngs.schedule_batch_noalloc(view, P, d_pc.data_ptr(), d_nc.data_ptr(), max_c, d_pol.data_ptr(), got.data_ptr(),
                                            got.data_ptr() + 4 * P, got.data_ptr() + 8 * P, load=load, seed=9, seq0=call * P)
            assert rc == ngs.OK, rc
            assert torch.equal(got, want), f"call {call}: decisions differ"
            if call == 0:  # the oracle itself on the first call (the snapshot is still the generated one)
                import oracle as o

                oc = o.OracleCluster(n)
                oc.load_snapshot(gcount, free)
                ol = None
                if load_on:
                    for m, (value, stamp, absent) in enumerate(load_tables):
                        oc.load_update_bulk(m, value, stamp, ok=np.repeat(~absent[:, None], 8, axis=1).astype(np.uint8))
                    ol = o.make_load_ctx(wl.DEFAULT_PERIODS, NOW_NS)
                onode, oscore, ogidx = oc.schedule_batch(pc, nc, pol, load=ol, seed=9, seq0=0)
                g = got.cpu().numpy()
                assert np.array_equal(g[:P], onode) and np.array_equal(g[P:2 * P], oscore)
                assert np.array_equal(g[2 * P:].reshape(P, max_c), ogidx)
                oc.close()
            assert (want[:P] >= 0).sum().item() > P // 2
            _, fr, rl = ref.snapshot_read()
            assert np.array_equal(t["free"].cpu().numpy(), fr), f"call {call}: snapshots differ"
            assert np.array_equal(t["remain_load"].cpu().numpy(), rl)


def test_noalloc_limits_and_arguments(built):
    import torch

    import nano_gpu_scheduler_b200 as ngs

    n, max_c = 64, 2
    dev = torch.device("cuda", 0)
    t = _tables_on_device(torch, np.full(n, 8, dtype=np.uint8), np.full((n, 8), 100, dtype=np.int32), None)
    view = _view(ngs, t, n)
    P = ngs._abi.NOALLOC_MAX_PODS + 1
    pc = torch.full((P, max_c), 10, dtype=torch.int32, device=dev)
    nc = torch.full((P,), 2, dtype=torch.int32, device=dev)
    pol = torch.zeros(P, dtype=torch.int32, device=dev)
    out = torch.zeros((2 + max_c) * P, dtype=torch.int32, device=dev)
    before = t["free"].clone()
    args = (d.data_ptr() for d in (pc, nc))
    rc = ngs.schedule_batch_noalloc(view, P, pc.data_ptr(), nc.data_ptr(), max_c, pol.data_ptr(), out.data_ptr(),
                                    out.data_ptr() + 4 * P, out.data_ptr() + 8 * P)
    assert rc == ngs._abi.ERR_NOALLOC_LIMIT  # one pod too many: refused, nothing launched
    assert b"NGS_NOALLOC_MAX_PODS" in ngs._abi.load().ngs_strerror(rc)
    assert torch.equal(t["free"], before)
    # the largest batch that fits is fine
    P = ngs._abi.NOALLOC_MAX_PODS
    rc = ngs.schedule_batch_noalloc(view, P, pc.data_ptr(), nc.data_ptr(), max_c, pol.data_ptr(), out.data_ptr(),
                                    out.data_ptr() + 4 * P, out.data_ptr() + 8 * P)
    assert rc == ngs.OK and (out[:P] >= 0).all()
    # a bad pod is reported by the batch's status word and by its own node_out; the others are scheduled
    pc[3, 0] = -7
    rc = ngs.schedule_batch_noalloc(view, P, pc.data_ptr(), nc.data_ptr(), max_c, pol.data_ptr(), out.data_ptr(),
                                    out.data_ptr() + 4 * P, out.data_ptr() + 8 * P)
    assert rc == ngs.ERR_INVALID
    assert out[3].item() == ngs._abi.NODE_INVALID_POD and (out[:3] >= 0).all() and (out[4:P] >= 0).all()
    # a table at an address the vector loads cannot take is refused before anything is launched
    odd = torch.zeros(n * 8 + 1, dtype=torch.int32, device=dev)
    bad_view = _view(ngs, dict(t, free=odd[1:]), n)
    rc = ngs.schedule_batch_noalloc(bad_view, 4, pc.data_ptr(), nc.data_ptr(), max_c, pol.data_ptr(), out.data_ptr(),
                                    out.data_ptr() + 16, out.data_ptr() + 32)
    assert rc == ngs.ERR_INVALID
    assert ngs.schedule_batch_noalloc(view, 1, 0, nc.data_ptr(), max_c, pol.data_ptr(), out.data_ptr(), out.data_ptr(),
                                      out.data_ptr()) == ngs.ERR_INVALID
