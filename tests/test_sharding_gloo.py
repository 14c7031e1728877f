"""The N>1 host path on CPU: two processes, gloo backend, timepoints sharded by block,
boundary node table exchanged between ranks (frames.send_frame / recv_frame).  The compute
calls go to the host-emulation harness; the result must equal the single-process database."""
import os
import shutil
import subprocess
import sys
import textwrap
from pathlib import Path

import pytest

from ultrack_b200 import sharding

ROOT = Path(__file__).resolve().parent.parent


def test_link_pairs_follow_segment_partition():
    for T in (2, 5, 7, 10, 50):
        for world in (2, 3, 4, 8):
            for r in range(world):
                seg = sharding.owned(range(T), T, world, r)
                lnk = sharding.owned(range(T - 1), T, world, r)
                assert lnk == [t for t in seg if t < T - 1]


def test_block_bounds_partition():
    for n in (0, 1, 5, 8, 50, 400):
        for world in (1, 2, 3, 8):
            got = []
            for r in range(world):
                lo, hi = sharding.block_bounds(n, world, r)
                got += list(range(lo, hi))
                for i in range(lo, hi):
                    assert sharding.owner_of(i, n, world) == r
            assert got == list(range(n))
    assert list(sharding.shard(range(10), 4, 1)) == [3, 4, 5]


def test_batch_index_range_matches_reference():
    assert list(sharding.batch_index_range(10, 4, None)) == list(range(10))
    assert list(sharding.batch_index_range(10, 4, 2)) == [8, 9]
    with pytest.raises(IndexError):
        sharding.batch_index_range(8, 4, 2)


WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, {root!r})
    import torch.distributed as dist
    from tests.emu_harness import emu_lib
    from ultrack_b200 import _lib, MainConfig, segment, link, synth, frames
    _lib._LIB = emu_lib()          # test harness: host emulation of the CUDA kernels
    dist.init_process_group("gloo")
    cfg = MainConfig.model_validate({{"data": {{"working_dir": {wd!r}}},
        "segmentation": {{"min_area": 60, "max_area": 2000}}, "linking": {{"max_distance": 10, "max_neighbors": 3}}}})
    fg, ct = synth.make_timelapse((72, 72), {T}, seed=5, variant="S", radius=6.0, motion=2.0)
    segment(fg, ct, cfg)
    if {drop_store}:
        frames.STORE.clear()       # link() then rebuilds tables from the database blobs
    else:
        # link pairs follow segment()'s partition: every table link() needs is resident or
        # arrives through the boundary exchange, never through the database blobs
        def _no_db(*a, **k):
            raise AssertionError("link() fell back to frame_from_database")
        frames.frame_from_database = _no_db
    link(cfg)
    dist.barrier()
    dist.destroy_process_group()
""")


def _run(tmp_path, world, T, drop_store=False):
    wd = tmp_path / f"w{world}_{T}_{int(drop_store)}"
    wd.mkdir()
    script = wd / "worker.py"
    script.write_text(WORKER.format(root=str(ROOT), wd=str(wd), T=T, drop_store=drop_store))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 2000 + world), str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    from ultrack_b200 import database
    with database.connect(f"sqlite:///{wd}/data.db") as conn:
        nodes = conn.execute("select t,id,t_hier_id,z,y,x,area,hier_parent_id,pickle from nodes order by id").fetchall()
        ov = sorted(conn.execute("select node_id, ancestor_id from overlaps").fetchall())
        links = sorted(conn.execute("select source_id,target_id,weight from links").fetchall())
    return nodes, ov, links


@pytest.mark.parametrize("world,T", [(2, 5), (2, 2), (3, 7)])
def test_ranks_equal_one(tmp_path, world, T):
    """T is not a multiple of the world size: segment and link must still agree on who owns what."""
    one = _run(tmp_path, 1, T)
    two = _run(tmp_path, world, T)
    assert one[0] == two[0] and len(one[0]) > 0
    assert one[1] == two[1]
    assert one[2] == two[2] and len(one[2]) > 0


def test_two_ranks_link_from_database(tmp_path):
    one = _run(tmp_path, 1, 4)
    two = _run(tmp_path, 2, 4, drop_store=True)
    assert one[2] == two[2] and len(one[2]) > 0
