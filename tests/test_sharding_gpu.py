"""The N>1 path on real GPUs: two ranks over NCCL (timepoint blocks, boundary node table sent
device-to-device) must write the same database as one rank.  Skipped with fewer than 2 GPUs."""
import os
import subprocess
import sys
import textwrap
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent

WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, {root!r})
    import torch, torch.distributed as dist
    from ultrack_b200 import MainConfig, segment, link, synth
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    cfg = MainConfig.model_validate({{"data": {{"working_dir": {wd!r}}},
        "segmentation": {{"min_area": 60, "max_area": 2000}}, "linking": {{"max_distance": 10, "max_neighbors": 3}}}})
    fg, ct = synth.make_timelapse((24, 64, 64), {T}, seed=5, variant="S", radius=6.0, motion=2.0)
    segment(fg, ct, cfg)
    link(cfg)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
""")


def _run(tmp_path, world, T):
    wd = tmp_path / f"g{world}_{T}"
    wd.mkdir()
    script = wd / "worker.py"
    script.write_text(WORKER.format(root=str(ROOT), wd=str(wd), T=T))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    if world == 1:
        cmd = [sys.executable, str(script)]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", str(29700 + os.getpid() % 200), str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    from ultrack_b200 import database
    with database.connect(f"sqlite:///{wd}/data.db") as conn:
        nodes = conn.execute("select t,id,t_hier_id,z,y,x,area,hier_parent_id,pickle from nodes order by id").fetchall()
        ov = sorted(conn.execute("select node_id, ancestor_id from overlaps").fetchall())
        links = sorted(conn.execute("select source_id,target_id,weight from links").fetchall())
    return nodes, ov, links


@pytest.mark.gpu
@pytest.mark.parametrize("world,T", [(2, 5), (4, 9), (8, 19)])
def test_ranks_equal_one_gpu(tmp_path, world, T):
    """`world` ranks over NCCL (T not a multiple of the world size: shards of unequal length, the
    boundary node table of every shard but the last received device-to-device) write the same
    database as one rank.  Skipped when the box has fewer GPUs."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    one = _run(tmp_path, 1, T)
    many = _run(tmp_path, world, T)
    assert one[0] == many[0] and len(one[0]) > 0
    assert one[1] == many[1]
    assert one[2] == many[2] and len(one[2]) > 0
    print(f"sharding parity: {world} NCCL ranks == 1 rank on T={T}: {len(one[0])} nodes, {len(one[2])} links")
