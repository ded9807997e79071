"""N > 1 host logic on CPU: world_size-2 gloo processes (no GPU).  The data path has no
collective (frames are sharded); what is distributed is the bookkeeping bench.py relies on:
disjoint covering shards, and max/sum over ranks for the whole-job throughput."""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, n_frames: int, out_dir: str):
    sys.path.insert(0, str(ROOT / "stereo-matching_b200"))
    # import the helper module directly: the package __init__ needs the CUDA library
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "stm_sharding", str(ROOT / "stereo-matching_b200" / "stereomatch" / "sharding.py"))
    sh = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sh)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b, e = sh.shard_range(n_frames, rank, world)
        mine = torch.zeros(n_frames, dtype=torch.int64)
        mine[b:e] = 1
        dist.all_reduce(mine)  # every frame owned exactly once
        assert bool((mine == 1).all()), mine.tolist()
        # timing bookkeeping: rank r "took" 10 + r ms and processed (e - b) frames
        assert sh.max_over_ranks(10.0 + rank) == 10.0 + (world - 1)
        assert sh.sum_over_ranks(e - b) == n_frames
        seeds = list(sh.frame_seeds(1000, 3, rank))
        assert seeds == [1000 + 3 * rank + i for i in range(3)]
        dist.barrier()
        Path(out_dir, f"ok{rank}").write_text(f"{b} {e}")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_frames", [64, 7, 1])
def test_frame_sharding_world_size_2_gloo(tmp_path, n_frames):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n_frames, str(tmp_path)), nprocs=world, join=True)
    spans = [tuple(map(int, (tmp_path / f"ok{r}").read_text().split())) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n_frames
    assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))


def test_shard_range_properties():
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "stm_sharding", str(ROOT / "stereo-matching_b200" / "stereomatch" / "sharding.py"))
    sh = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sh)
    for n in (0, 1, 5, 64, 65):
        for w in (1, 2, 3, 8):
            spans = [sh.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sh.shard_range(4, 2, 2)
    assert sh.max_over_ranks(3.5) == 3.5 and sh.sum_over_ranks(2) == 2.0  # not distributed
