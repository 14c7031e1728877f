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
def test_two_gpus_equal_one(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    one = _run(tmp_path, 1, 5)
    two = _run(tmp_path, 2, 5)
    assert one[0] == two[0] and len(one[0]) > 0
    assert one[1] == two[1]
    assert one[2] == two[2] and len(one[2]) > 0
