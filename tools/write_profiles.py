"""Turns the raw files of a measurement round (gpurun_out/<tag>_bench_{1,2,4,8}gpu.json, <tag>_bench_reference.json,
<tag>_launches.csv) into the summaries under profiles/.  python tools/write_profiles.py r01h"""
import collections
import csv
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    tag = sys.argv[1]
    src, dst = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
    vals = {}
    for n in (1, 2, 4, 8):
        name = "%s_bench_%dgpu.json" % (tag, n)
        shutil.copy(os.path.join(src, name), os.path.join(dst, name))
        vals[n] = json.load(open(os.path.join(dst, name)))
    shutil.copy(os.path.join(src, tag + "_bench_reference.json"), dst)
    shutil.copy(os.path.join(src, tag + "_launches.csv"), dst)
    ref = json.load(open(os.path.join(dst, tag + "_bench_reference.json")))
    with open(os.path.join(dst, tag + "_scaling.md"), "w") as f:
        f.write("# Round 1, final build: weak scaling, BASELINE configs[1] (8 images per GPU per step)\n\n")
        f.write("Launched as the driver does: `python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr "
                "127.0.0.1 bench.py --gpus N --steps 200 --warmup 20` (N = 1: plain `python bench.py`).  One box per N; timing = "
                "CUDA events, max over ranks.  Raw lines: `%s_bench_{1,2,4,8}gpu.json`.\n\n" % tag)
        f.write("| GPUs | images/s (device-resident) | ms per step | x of 1 GPU | e2e images/s (host buffers, pipelined) | "
                "x of 1 GPU | SM clock MHz |\n|---|---|---|---|---|---|---|\n")
        for n in (1, 2, 4, 8):
            d = vals[n]
            f.write("| %d | %.0f | %.4f | %.2f | %.0f | %.2f | %s |\n" % (
                n, d["value"], d["ms_per_step"], d["value"] / vals[1]["value"], d["e2e"]["value"],
                d["e2e"]["value"] / vals[1]["e2e"]["value"], d["clocks"]["sm_mhz"]))
        d = vals[1]
        f.write("\n1 GPU detail: one launch per step (cluster size chosen by the library: 10 CTAs per image at this batch), "
                "phases (us) %s; saturating batch (256 images per call) %.0f images/s; two batches in flight %.0f images/s; "
                "e2e H2D rate %.1f GB/s against %.1f GB/s for a plain pinned copy of the same bytes (the host link is the bound; "
                "with several ranks the host's memory and PCIe complex are shared).\n" % (
                    json.dumps({k: round(v, 2) for k, v in d["fused_kernel_phases_us"].items()}),
                    d["saturating_batch"]["value"], d["two_batches_in_flight"]["value"], d["e2e"]["h2d_gbs"],
                    d["e2e"]["h2d_link_gbs_plain_copy"]))
        f.write("\nReference arm on the same pool (`bench.py --impl reference --steps 3 --warmup 1`, numpy restatement, %d worker "
                "processes): %.1f images/s.  CPU baseline inside the 1-GPU line: %.1f images/s on %d cores.\n" % (
                    ref["cpu_baseline"]["cores"], ref["value"], d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"]))
    lines = [ln for ln in open(os.path.join(dst, tag + "_launches.csv")) if not ln.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        k = (row["Kernel Name"], row["Grid Size"], row["Block Size"])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(row["Metric Value"].replace(",", ""))
    with open(os.path.join(dst, tag + "_launches_summary.md"), "w") as f:
        f.write("# Round 1, final build: ncu launch list\n\nCommand: `ncu --metrics gpu__time_duration.sum --clock-control none -c 600 "
                "--csv python bench.py --steps 20 --warmup 3`\n(the same command exited 0 without ncu first; B200; per-launch times "
                "are cold-cache and serialised: compare SHARES).  Raw list: `%s_launches.csv`.\n\n" % tag)
        f.write("| kernel | grid | block | launches | mean us/launch |\n|---|---|---|---|---|\n")
        for (k, g, b), (n, t) in agg.items():
            f.write("| %s | %s | %s | %d | %.1f |\n" % (k.replace("void ", "").replace("|", "/")[:110], g, b, n, t / n / 1e3))
        f.write("\nOne proposal-op call (batch 8) is ONE launch of `proposal_cluster_kernel<4>`: grid 80 = 8 clusters x 10 CTAs (the "
                "library picks the cluster size per call), 1024 threads: 100% of the step.  The two-batches-in-flight figure launches "
                "the same grid from a second handle; the grid-512 rows of `proposal_cluster_kernel<12>` are bench.py's saturating batch "
                "(256 images as 2-CTA clusters, 12 keys per thread); `alu_stream_kernel` is the ALU-peak micro-benchmark; the torch "
                "kernels are harness fills / copies / the result check.\n")
    print(open(os.path.join(dst, tag + "_scaling.md")).read())


if __name__ == "__main__":
    main()
