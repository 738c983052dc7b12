"""NVLink result gather: decode the lineitem-shaped SF10 shard on GPU 0 and copy all 16 decoded columns into buffers on
GPU 1 with flsgpu_batch_copy_to_device (optional).  Prints one JSON line: bytes, best time of
`passes` (sum of the per-column device times), GB/s.  Needs 2 GPUs; with one it copies device 0 -> device 0.
Usage: python profiles/gather_bench.py [rowgroups] [passes]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "duckdb-fastlane_b200"))
sys.path.insert(0, ROOT)
import __graft_entry__  # noqa: E402

__graft_entry__.build()
import torch  # noqa: E402
import flsgpu  # noqa: E402
import synth  # noqa: E402
import fls_writer as W  # noqa: E402

n_rg = int(sys.argv[1]) if len(sys.argv) > 1 else 916
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dst = 1 if torch.cuda.device_count() > 1 else 0
img = W.tile_rowgroups(synth.lineitem_image(64, seed=7), n_rg)
with flsgpu.Reader.from_image(img, device=0) as rd:
    rd.make_resident()
    with rd.decode_batch(0, rd.n_rowgroups) as b:
        bufs, total = [], 0
        for i in range(b.n):
            c = b.columns[i]
            if c.width:
                v = torch.empty(c.padded_rows * c.width, dtype=torch.uint8, device="cuda:%d" % dst)
                bufs.append((v, None))
                total += v.numel()
            else:
                v = torch.empty(max(int(c.payload_bytes), 1), dtype=torch.uint8, device="cuda:%d" % dst)
                o = torch.empty(c.padded_rows + 1, dtype=torch.int32, device="cuda:%d" % dst)
                bufs.append((v, o))
                total += int(c.payload_bytes) + 4 * o.numel()
        best = None
        for _ in range(passes + 1):
            ms = 0.0
            for i, (v, o) in enumerate(bufs):
                ms += b.copy_to_device(i, dst, v.data_ptr(), o.data_ptr() if o is not None else None)
            best = ms if best is None else min(best, ms)
        can = torch.cuda.can_device_access_peer(0, dst) if dst else False
print(json.dumps({"src_device": 0, "dst_device": dst, "peer": bool(can), "bytes": total, "ms_best": round(best, 3),
                  "GBps": round(total / best / 1e6, 1), "columns": len(bufs)}))
