"""CPU tests of the host logic and the C-ABI surface (no compute calls: no GPU here)."""
import ctypes
import itertools
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import acoss_b200
    from acoss_b200 import _lib
    L = acoss_b200.load_library()
    hdr = open(os.path.join(ROOT, "include", "acoss_b200.h")).read()
    declared = set(re.findall(r"\b(acoss_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(L, name), "missing export %s" % name
    assert declared == set(_lib.SYMBOLS), (declared ^ set(_lib.SYMBOLS))
    assert b"sm_100a" in L.acoss_version()


def test_struct_layouts_match_header():
    from acoss_b200 import _lib
    assert ctypes.sizeof(_lib.CrpParams) == 28
    assert ctypes.sizeof(_lib.PairStats) == 4 * 8 + 10 * 8


def test_no_cpu_fallback_without_gpu():
    import torch
    import acoss_b200
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(acoss_b200.AcossError):
        acoss_b200.Context(0)


def test_np_pairwise_sum_matches_numpy():
    import acoss_b200
    L = acoss_b200.load_library()
    rng = np.random.default_rng(0)
    for n in [0, 1, 2, 7, 8, 9, 12, 15, 16, 17, 100, 127, 128, 129, 255, 1000, 1001, 4097, 15000]:
        a = (rng.random(n) * 10).astype(np.float64)
        got = L.acoss_np_pairwise_sum(a.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), n)
        assert got == float(np.sum(a)), n
        if n:
            assert got / n == float(np.mean(a)), n


def test_enumerate_pairs_matches_itertools():
    from acoss_b200 import CoverAlgorithm
    for n in (2, 3, 7, 20):
        assert CoverAlgorithm.enumerate_pairs(n, True).tolist() == [list(p) for p in itertools.combinations(range(n), 2)]
        assert CoverAlgorithm.enumerate_pairs(n, False).tolist() == [list(p) for p in itertools.permutations(range(n), 2)]


def test_shard_pairs_partition_and_balance():
    from acoss_b200.algorithm_template import shard_pairs
    rng = np.random.default_rng(1)
    pairs = np.stack(np.triu_indices(60, 1), 1)
    cost = rng.integers(1, 1000, size=len(pairs))
    for world in (2, 3, 8):
        parts = [shard_pairs(pairs, cost, r, world) for r in range(world)]
        allidx = np.sort(np.concatenate(parts))
        assert np.array_equal(allidx, np.arange(len(pairs)))
        loads = np.array([cost[p].sum() for p in parts], dtype=float)
        assert loads.max() / loads.mean() < 1.05


def test_create_dataset_filepaths_and_errors(tmp_path):
    from acoss_b200.utils import create_dataset_filepaths
    p = tmp_path / "a.csv"
    p.write_text("work_id,track_id\nW_1,P_1\nW_1,P_2\n")
    assert create_dataset_filepaths(str(p), "feat/", ".h5") == ["feat/W_1/P_1.h5", "feat/W_1/P_2.h5"]
    q = tmp_path / "b.csv"
    q.write_text("work,track_id\nW_1,P_1\n")
    with pytest.raises(IOError):
        create_dataset_filepaths(str(q), "feat/", ".h5")


def test_benchmark_unknown_algorithm_raises(workdir):
    import acoss_b200
    with pytest.raises(NotImplementedError):
        acoss_b200.benchmark("x.csv", "feat/", algorithm="Nope")


def test_synth_dataset_roundtrip_and_clique_info(workdir):
    import acoss_b200
    from acoss_b200 import synth
    ds = synth.make_named("covers80", "S", seed=1, scale=0.05)
    fdir = synth.write_dataset(str(workdir / "feat"), ds, str(workdir / "ds.csv"))
    lens = [h.shape[0] for h in ds["hpcp"]]
    assert len(set(lens)) == len(lens)
    algo = acoss_b200.CoverAlgorithm(str(workdir / "ds.csv"), datapath=fdir, shortname="t")
    assert algo.N == len(lens) and algo.Ds["main"].shape == (algo.N, algo.N)
    f = algo.load_features(1)
    assert f["label"] == ds["labels"][1] and np.array_equal(f["hpcp"], ds["hpcp"][1])
    algo.get_all_clique_ids()
    assert sum(len(c) for c in algo.cliques.values()) == algo.N
    order, csize, cstart = algo.clique_layout()
    assert sorted(order.tolist()) == list(range(algo.N)) and (csize == 2).all()
    algo.cleanup_memmap()
    assert not os.path.exists("%s_main_dmat" % algo.get_cacheprefix())


GLOO_SCRIPT = r"""
import os, sys
import numpy as np
sys.path.insert(0, %r)
import torch.distributed as dist
from acoss_b200.algorithm_template import shard_pairs, all_gather_scores, dist_info
dist.init_process_group("gloo")
rank, world = dist_info()
pairs = np.stack(np.triu_indices(23, 1), 1)
cost = (pairs[:, 0] + 1) * (pairs[:, 1] + 3)
mine = shard_pairs(pairs, cost, rank, world)
truth = np.stack([pairs[:, 0] * 100.0 + pairs[:, 1], pairs[:, 0] - 0.5 * pairs[:, 1]]).astype(np.float32)
full = all_gather_scores(mine, truth[:, mine], len(pairs))
assert full.shape == truth.shape and np.array_equal(full, truth), "gather mismatch on rank %%d" %% rank
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_gloo_world2_shard_and_allgather(tmp_path):
    script = tmp_path / "w2.py"
    script.write_text(GLOO_SCRIPT % ROOT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29531", str(script)],
                       capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("ok") == 2


def _simple_ftm_dataset(workdir):
    from acoss_b200 import synth
    from oracle.make_golden import simple_ftm_dataset
    ds, extra = simple_ftm_dataset()
    fdir = synth.write_dataset(str(workdir / "feat"), ds, str(workdir / "ds.csv"), extra=extra)
    return str(workdir / "ds.csv"), fdir, ds


def test_feature_ports_match_reference_goldens(golden, workdir):
    """oracle/ports.py restatements of the per-track SiMPle / FTM2D features against the reference run verbatim (the
    product classes compute them on the device: tests/test_gpu_parity.py)."""
    from oracle import ports as P
    from oracle.make_golden import simple_ftm_dataset
    g = golden("simple_ftm2d_reference.npz")
    ds, extra = simple_ftm_dataset()
    assert np.allclose([float(np.sum(h, dtype=np.float64)) for h in ds["hpcp"]], g["hpcp_sum"], rtol=0, atol=0)
    for i, h in enumerate(ds["hpcp"]):
        np.testing.assert_allclose(P.simple_features(h.T), g["simple_feat%d" % i], rtol=1e-12, atol=1e-15)
        np.testing.assert_allclose(P.ftm2d_shingle(h.T, extra["madmom_features"][i]["onsets"]), g["ftm_shingle%d" % i], rtol=1e-10, atol=1e-14)


def test_packed_feature_store_roundtrip_and_errors(tmp_path):
    """featio.save_packed / load_packed: ragged tracks incl. an empty one, labels, mmap view,
    and loud failures on foreign / truncated / inconsistent files."""
    from acoss_b200 import featio
    rng = np.random.default_rng(3)
    tracks = [rng.random((n, 12)).astype(np.float32) for n in (17, 0, 1, 2049, 300)]
    labels = ["w0", "w0", "wörk 2", "w3", "w3"]
    p = str(tmp_path / "x.acosspk")
    featio.save_packed(p, tracks, labels)
    frames, offs, lab = featio.load_packed(p)
    assert isinstance(frames, np.memmap) and frames.dtype == np.float32 and frames.shape == (sum(t.shape[0] for t in tracks), 12)
    assert lab == labels and offs.tolist() == [0, 17, 17, 18, 2067, 2367]
    for k, t in enumerate(tracks):
        assert np.array_equal(np.asarray(frames[offs[k]:offs[k + 1]]), t)
    assert frames.offset % 64 == 0                              # data section is 64-byte aligned
    with pytest.raises(ValueError):
        featio.save_packed(p + "2", tracks, labels[:-1])
    with pytest.raises(ValueError):
        featio.save_packed(p + "2", [tracks[0], tracks[0][:, :5]], ["a", "b"])
    with pytest.raises(ValueError):
        featio.save_packed(p + "2", tracks[:1], ["bad\nlabel"])
    bad = str(tmp_path / "bad.acosspk")
    open(bad, "wb").write(b"not a store" * 20)
    with pytest.raises(IOError):
        featio.load_packed(bad)
    raw = bytearray(open(p, "rb").read())
    raw[64 + 8:64 + 16] = np.array([-5], np.int64).tobytes()    # offsets no longer monotone
    open(bad, "wb").write(bytes(raw))
    with pytest.raises(IOError):
        featio.load_packed(bad)
    open(bad, "wb").write(bytes(open(p, "rb").read()[:40]))     # truncated header
    with pytest.raises(IOError):
        featio.load_packed(bad)
