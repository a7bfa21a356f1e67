#!/usr/bin/env python
"""profiles/r02_configs_scaling.json (+ a markdown table on stdout) from the bench lines profiles/r02_bench_n{1,2,4,8}.jsonl
(`bench.py --config all` alone and under torchrun).  Usage: python tools/scaling_table.py [prefix=profiles/r02_bench_n]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def lines(path):
    if not os.path.exists(path):
        return []
    return [json.loads(l) for l in open(path) if l.startswith("{")]


def main(prefix):
    table = {"source": "bench.py --config all at N=1 and under torchrun at N=2,4,8 on one 8xB200 node; value = device-timed resident Msamples/s, max over ranks; "
                       "e2e = ONE process, C-ABI multi-device context, host buffers, unchanged scene between steps (N=1: renderer.render); "
                       "first_call_ms = the e2e call that copied the scene and built its structures; e2e_torchrun = one process per GPU; "
                       "cpu_baseline = the unmodified reference on the box's host threads", "configs": {}}
    for n in (1, 2, 4, 8):
        for d in lines(os.path.join(ROOT, f"{prefix}{n}.jsonl")):
            c = table["configs"].setdefault(d["config"]["name"], {"metric": d["metric"], "n": {}})
            e = {"value": d["value"], "mrays_per_s": d["mrays_per_s"], "ms_per_step": d["ms_per_step"], "e2e": d["e2e"]["value"],
                 "e2e_ms": d["e2e"]["ms_per_step"], "e2e_first_call_ms": d["e2e"].get("first_call_ms"), "clocks": d["clocks"]}
            if "ms_reduce" in d["e2e"]:
                e["nccl_reduce_ms"] = d["e2e"]["ms_reduce"]
            if "e2e_torchrun" in d:
                e["e2e_torchrun"] = d["e2e_torchrun"]["value"]
            if "cpu_baseline" in d:
                c["cpu_baseline"] = d["cpu_baseline"]
            if n == 1:
                c["roofline_frac"], c["roofline_kernel"] = d["roofline"]["frac"], d["roofline"]["kernel"]
            c["n"][str(n)] = e
    print("| config | CPU | N=1 value / e2e | N=2 | N=4 | N=8 | e2e C ABI 2/4/8 | e2e torchrun 2/4/8 | ncclReduce ms 2/4/8 |")
    print("|---|---|---|---|---|---|---|---|---|")
    for name, c in table["configs"].items():
        if "1" not in c["n"]:
            continue
        v1 = c["n"]["1"]["value"]
        for n, e in c["n"].items():
            e["efficiency_value"] = e["value"] / (v1 * int(n))

        def g(n, k, fmt="{:.0f}"):
            e = c["n"].get(str(n))
            return fmt.format(e[k]) if e and e.get(k) is not None else "-"
        cpu = c.get("cpu_baseline", {}).get("value")
        print(f"| {name} | {cpu:.3g} | {g(1, 'value')} / {g(1, 'e2e')} | " +
              " | ".join(f"{g(n, 'value')} ({g(n, 'efficiency_value', '{:.3f}')})" for n in (2, 4, 8)) + " | " +
              " / ".join(g(n, "e2e") for n in (2, 4, 8)) + " | " + " / ".join(g(n, "e2e_torchrun") for n in (2, 4, 8)) + " | " +
              " / ".join(g(n, "nccl_reduce_ms", "{:.2f}") for n in (2, 4, 8)) + " |")
    with open(os.path.join(ROOT, "profiles", "r02_configs_scaling.json"), "w") as f:
        json.dump(table, f, indent=1)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "profiles/r02_bench_n")
