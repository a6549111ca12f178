"""N>1 host logic on CPU: two gloo ranks shard a frame list, "compress" their ranges (sizes only) and agree on the
EBCK container layout without any collective on frame data."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys, json
    sys.path.insert(0, %r)
    import torch.distributed as dist
    from ebcc_b200 import sharding as S
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%%s" %% os.environ["PORT"], rank=int(os.environ["RANK"]), world_size=2)
    F = 11
    lo, hi = S.frame_range(F, dist.get_rank(), 2)
    local = [1000 + 7 * t for t in range(lo, hi)]          # stand-in for per-frame stream sizes
    sizes = S.gather_sizes(local)
    offs, total = S.container_offsets(sizes)
    print(json.dumps({"rank": dist.get_rank(), "range": [lo, hi], "sizes": sizes, "offs": offs, "total": total}))
    dist.destroy_process_group()
""") % ROOT


def test_two_rank_sharding_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(WORKER)
    import json, socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(os.environ, RANK=str(r), PORT=str(port)),
                              stdout=subprocess.PIPE, text=True) for r in range(2)]
    outs = [json.loads(p.communicate(timeout=120)[0].strip().splitlines()[-1]) for p in procs]
    assert all(p.returncode == 0 for p in procs)
    a, b = sorted(outs, key=lambda o: o["rank"])
    assert a["range"] == [0, 6] and b["range"] == [6, 11]
    want = [1000 + 7 * t for t in range(11)]
    assert a["sizes"] == want and b["sizes"] == want and a["offs"] == b["offs"]
    assert a["offs"][0] == 80 and a["offs"][1] == 80 + 8 + 1000 and a["total"] == 80 + sum(8 + s for s in want)
