"""SQL-level throughput of the table function (SURVEY.md section 8d last row): a DuckDB v1.3.2 instance with read_fls
registered (ext/duckdb_sql_check), SET threads = 1 and = nproc, against the same query over the same table held in
DuckDB's own storage (CREATE TABLE .. AS SELECT * FROM read_fls).  One JSON line per measurement.
Usage: python profiles/sql_bench.py [rowgroups=92 (SF1)] [repeats=3]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "duckdb-fastlane_b200"))
sys.path.insert(0, ROOT)
import synth  # noqa: E402
import fls_writer as W  # noqa: E402

EXT = os.path.join(ROOT, "duckdb-fastlane_b200", "ext")
n_rg = int(sys.argv[1]) if len(sys.argv) > 1 else 92
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
small = synth.lineitem_image(min(64, n_rg), seed=7)
img = W.tile_rowgroups(small, n_rg) if n_rg > 64 else small
path = "/dev/shm/flsgpu_sql_%d.fls" % os.getpid()
img.tofile(path)
rows = n_rg * 65536
dec_bytes = None
env = dict(os.environ)
env["LD_LIBRARY_PATH"] = os.path.join(EXT, "_duckdb") + ":" + os.path.join(ROOT, "duckdb-fastlane_b200") + ":" + env.get("LD_LIBRARY_PATH", "")
ncpu = os.cpu_count() or 1
queries = {
    "agg3": "SELECT count(*), sum(l_extendedprice), max(length(l_comment)) FROM %s",
    "count": "SELECT count(*) FROM %s",
    "sum_price": "SELECT sum(l_extendedprice) FROM %s",
    "q6": "SELECT sum(l_extendedprice * l_discount) FROM %s WHERE l_shipdate >= DATE '1994-01-01' AND l_shipdate < DATE '1995-01-01' "
          "AND l_discount BETWEEN 0.05 AND 0.07 AND l_quantity < 24",
    "shipdate_1pct": "SELECT count(*), sum(l_extendedprice), max(l_comment) FROM %s WHERE l_shipdate >= DATE '1995-06-01' AND l_shipdate < DATE '1995-06-26'",
}
try:
    for threads in (1, ncpu):
        stm = ["SET threads = %d" % threads, "CREATE TABLE li AS SELECT * FROM read_fls('%s')" % path]
        tags = [None, ("load", "create table as select")]
        for name, q in queries.items():
            for src, tag in (("read_fls('%s')" % path, "read_fls"), ("li", "duckdb_table")):
                for _ in range(reps):
                    stm.append(q % src)
                    tags.append((name, tag))
        r = subprocess.run([os.path.join(EXT, "duckdb_sql_check")] + stm, capture_output=True, text=True, env=env, timeout=3000)
        out = [json.loads(line) for line in r.stdout.strip().splitlines() if line.startswith("{")]
        best = {}
        for o, t in zip(out, tags):
            if t is None:
                continue
            if "error" in o:
                print(json.dumps({"error": o["error"], "sql": o["sql"]}), flush=True)
                continue
            k = (t[0], t[1])
            if k not in best or o["ms"] < best[k][0]:
                best[k] = (o["ms"], o["rows"])
        for (name, tag), (ms, res) in best.items():
            print(json.dumps({"threads": threads, "query": name, "source": tag, "ms_best_of_%d" % reps: round(ms, 2),
                              "rows_scanned": rows, "Mrows_per_s": round(rows / ms / 1e3, 1), "result": res[0] if res else None}), flush=True)
finally:
    os.unlink(path)
