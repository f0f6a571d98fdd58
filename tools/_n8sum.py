import json
for f in ("p2p", "nccl"):
    try:
        d = json.load(open(f"gpurun_out/bench_n8_{f}.json"))
        print(f, d["value"], d["ms_per_step"], d["roofline"].get("k6_ms"), d["e2e"]["value"], d["config"].get("parallelism"))
        if "ensemble" in d:
            print("   ensemble", d["ensemble"]["value"], d["ensemble"]["ms_per_launch"])
    except Exception as e:
        print(f, "unreadable", e)
