"""Print one training step of a chrome trace written by `bench/ingest_bench.py --timeline` as a table
(start, duration, stream, kernel), plus the step periods -- the listings in profiles/ingest_timeline.md.

    python bench/timeline_table.py trace.json [step_index]
Streams are named by what runs on them (train = the graph replay, copy = host->device copies, pack = ingest
kernels, comm = all-reduce kernels)."""
import json
import sys


def main():
    path = sys.argv[1]
    step = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    ev = json.load(open(path))["traceEvents"]
    ks = sorted([e for e in ev if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")],
                key=lambda e: e["ts"])
    names = {}
    for e in ks:
        s = e["args"].get("stream")
        n = e["name"]
        kind = ("copy" if e["cat"] == "gpu_memcpy" and "HtoD" in n else "pack" if "ingest_" in n else
                "comm" if "allreduce" in n else "train" if "gemm_tn_fast" in n or "mlp_head" in n else None)
        if kind and (s not in names or (kind == "train" and names[s] != "comm")):
            names[s] = kind
    pubs = [i for i, e in enumerate(ks) if "publish_host" in e["name"]]
    a, b = pubs[step], pubs[step + 1]
    t0 = ks[a]["ts"]
    rows, prev = [], None
    for e in ks[a:b + 1]:
        st = names.get(e["args"].get("stream"), str(e["args"].get("stream")))
        nm = e["name"].split("(")[0].replace("rdp::", "").replace("void ", "")
        if e["cat"] == "gpu_memcpy":
            nm = "H2D copy" if "HtoD" in e["name"] else e["name"]
        if nm == "H2D copy" and prev and prev[2] == "H2D copy" and prev[1] == st:
            prev[4] += 1
            continue
        prev = [e["ts"] - t0, st, nm[:48], e["dur"], 0]
        rows.append(prev)
    for r in rows:
        extra = f"   (+{r[4]} more, back to back)" if r[4] else ""
        print(f"{r[0]:8.1f} us {r[3]:6.1f} us  {r[1]:5s} {r[2]}{extra}")
    periods = [ks[pubs[i + 1]]["ts"] - ks[pubs[i]]["ts"] for i in range(len(pubs) - 1)]
    print("step periods under the profiler (us):", [round(p, 1) for p in periods])


if __name__ == "__main__":
    main()
