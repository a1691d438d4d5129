"""CPU: host-side logic - synthetic generators (product copy == oracle copy), sharding, the cross-rank arg-min
reduction over gloo with world_size 2, hand-tree flattening."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_synthetic_generators_match_oracle_copies():
    from compliancedex_b200 import load_hand
    from compliancedex_b200 import synthetic as ps
    from oracle import gpis_oracle as og
    from oracle import pregrasp_oracle as op
    for a, b in zip(ps.synthetic_training_set(300, seed=4), og.synthetic_training_set(300, seed=4)):
        assert torch.equal(a, b)
    doc = load_hand("allegro")
    A = ps.synthetic_candidates(doc, 13, [0.01, -0.02, 0.03], seed=9)
    B = op.synthetic_candidates(doc, 13, [0.01, -0.02, 0.03], seed=9)
    for a, b in zip(A[:3], B[:3]):
        assert torch.equal(a, b)
    assert np.array_equal(A[3], B[3])


def test_query_stream_is_world_size_invariant():
    from compliancedex_b200.distributed import shard_range
    from compliancedex_b200.synthetic import synthetic_queries
    total = (1 << 20) + 12345
    full = synthetic_queries(total, 0, total)
    for world in (2, 3, 8):
        parts = [synthetic_queries(total, *shard_range(total, r, world)) for r in range(world)]
        assert torch.equal(torch.cat(parts), full)
    assert shard_range(10, 3, 4) == (9, 10) and shard_range(10, 0, 1) == (0, 10) and shard_range(3, 3, 4) == (3, 3)


def test_reduce_pairs_tie_break_and_nan():
    from compliancedex_b200.distributed import reduce_pairs
    assert reduce_pairs([(3.0, 7), (1.0, 9), (1.0, 4), (2.0, 0)]) == (1.0, 4)
    assert reduce_pairs([(float("nan"), 1), (5.0, 8)]) == (5.0, 8)
    assert reduce_pairs([(float("inf"), 3), (float("inf"), 0)]) == (float("inf"), 0)     # torch.argmin of all-inf -> 0
    assert reduce_pairs([(float("nan"), -1)]) == (float("inf"), -1)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _FakeOpt:
    """stands in for a ProbabilisticGraspOptimizer shard: argmin_into writes this rank's (best value, global index)"""

    def __init__(self, values, offset):
        self.device = torch.device("cpu")
        self.values, self.offset = values, offset

    def argmin_into(self, out):
        i = int(torch.argmin(self.values))
        out[0] = self.values[i]
        out[1] = float(self.offset + i)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from compliancedex_b200.distributed import global_argmin, shard_range
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    allv = torch.tensor([5.0, 2.0, 9.0, 2.0, 7.0, 1.5, 1.5, 8.0], dtype=torch.float64)
    lo, hi = shard_range(allv.numel(), rank, world)
    got = global_argmin(_FakeOpt(allv[lo:hi], lo), world)
    q.put((rank, got))
    dist.destroy_process_group()


def test_global_argmin_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0] == res[1] == (1.5, 5)        # same answer on every rank; first occurrence of the minimum


def test_hand_tree_flattening():
    from compliancedex_b200.hand import load_hand, tree_for
    for hand, nb in (("leap", 17), ("allegro", 22)):
        doc = load_hand(hand)
        t = tree_for(doc, doc["ee_link_name"], doc["ee_link_offset"])
        assert t.n_bodies == nb and t.n_dofs == 16
        assert sorted(t.dof[t.dof >= 0].tolist()) == list(range(16))       # dof i = i-th non-fixed link in link order
        assert (t.parent[1:] < np.arange(1, nb)).all()
        R = t.rot.reshape(nb, 3, 3).astype(np.float64)
        assert np.abs(R @ R.transpose(0, 2, 1) - np.eye(3)).max() < 1e-6
    leap = tree_for(load_hand("leap"), load_hand("leap")["ee_link_name"], load_hand("leap")["ee_link_offset"])
    assert (leap.sign[1:] == -1).all() and (leap.axis[1:] == 2).all()      # all Leap axes are 0 0 -1
    with pytest.raises(ValueError):
        tree_for(load_hand("leap"), ["fingertip"], [[0, 0, 0]])


@pytest.mark.parametrize("tag", ["default", "box018"])
def test_completion_assembly_matches_reference_script(tag, golden_dir):
    """assemble_training_set vs the reference script's own statements (train_gpis_completion.py:24-58 executed
    verbatim by oracle/make_golden_completion.py) - bit for bit."""
    from compliancedex_b200.completion import assemble_training_set, exterior_points
    g = np.load(os.path.join(golden_dir, "completion_coffeebottle.npz"))
    num_obs, num_gen, obs_noise, gen_noise, num_internals, boundary, bias = g[f"{tag}_args"]
    X1, y1, noise = assemble_training_set(g[f"{tag}_observed"], g[f"{tag}_completed"], int(num_obs), int(num_gen), float(obs_noise),
                                          float(gen_noise), int(num_internals), float(boundary), weights=g[f"{tag}_draws"])
    assert np.array_equal(X1.numpy(), g[f"{tag}_X1"]) and np.array_equal(y1.numpy(), g[f"{tag}_y1"])
    assert np.array_equal(noise.numpy(), g[f"{tag}_noise"])
    assert X1.dtype == torch.float64 and tuple(y1.shape) == (X1.shape[0], 1) and exterior_points(boundary).shape == (14, 3)
    # seeded generator path: same draws as torch.manual_seed(7); torch.rand
    gen = torch.Generator().manual_seed(7)
    X1b, _, _ = assemble_training_set(g[f"{tag}_observed"], g[f"{tag}_completed"], int(num_obs), int(num_gen), float(obs_noise),
                                      float(gen_noise), int(num_internals), float(boundary), generator=gen)
    assert np.array_equal(X1b.numpy(), g[f"{tag}_X1"])


def test_driver_scripts_reference_only_defined_globals():
    """bench.py / __graft_entry__.py run unattended on the GPU box: a name that no scope defines is a NameError there, minutes into the
    run.  Static check over every function scope: each implicitly-global name must be bound at module level or be a builtin."""
    import builtins
    import glob
    import os
    import symtable

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    files = (glob.glob(os.path.join(root, "bench*.py")) + [os.path.join(root, "__graft_entry__.py")]
             + glob.glob(os.path.join(root, "compliancedex_b200", "*.py")) + glob.glob(os.path.join(root, "scripts", "*.py"))
             + glob.glob(os.path.join(root, "oracle", "*.py")))
    assert len(files) > 20
    bad = []
    for path in files:
        src = open(path).read()
        top = symtable.symtable(src, path, "exec")
        module_names = {s.get_name() for s in top.get_symbols() if s.is_assigned() or s.is_imported() or s.is_namespace()}
        star_import = "import *" in src

        def walk(scope):
            for child in scope.get_children():
                for s in child.get_symbols():
                    n = s.get_name()
                    if s.is_referenced() and s.is_global() and not s.is_assigned() and not s.is_imported():
                        if n not in module_names and not hasattr(builtins, n) and not star_import and n not in ("__file__", "__name__", "__doc__"):
                            bad.append((os.path.relpath(path, root), child.get_name(), n))
                walk(child)
        walk(top)
    assert not bad, bad
