#!/usr/bin/env python
"""Runs the reference's own MoE-path scripts UNMODIFIED against the drop-in `sgl_kernel` package and turns
their printed verdicts into pass / fail (the scripts never assert, they print
"Comparing:  True/False ..." through utils.compare, /root/reference/utils.py:10-25, and are launched as
`python3 <file>`, /root/reference/run_bench_cpu.sh:20).

Each script is started exactly that way -- `python3 <file>` in its own process, cwd = the reference
directory -- with sgl_cpu_tests_b200/dropin first on PYTHONPATH so that `import sgl_kernel`,
`sgl_kernel.common_ops` and `sglang.srt.layers.amx_utils` resolve to this build.  Nothing of the reference
is imported, patched or copied into the repository; the directory is found through --ref / $SGLB_REFERENCE,
else ./baseline/_ref (git-ignored scratch that travels to the GPU box), else /root/reference.

  python tools/run_reference_scripts.py [--ref DIR] [--out profiles/r2_reference_scripts.txt] [--bench]

Exit status 1 if any verdict is False, a script crashes, or a script prints no verdict at all.
"""
import argparse
import os
import re
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "sgl_cpu_tests_b200", "dropin")

# script -> call sites it covers ("What's missing" 1)
TESTS = [
    ("test_moe.py", "bf16 fused_experts + grouped_topk_cpu, 13-arg vintage (:109)"),
    ("test_moe_int8.py", "int8 w8a8 fused_experts (:139-141)"),
    ("test_moe_fp8_ext.py", "fp8 w8a16 shared_expert (:65-67) and fused_experts (:122-124)"),
    ("test_moe_fp8.py", "fp8 block-quantised shared_expert (:91-92)"),
    ("test_shared_experts.py", "shared_expert bf16 / int8, 12-arg vintage (:85-87)"),
    ("test_grouped_topk.py", "grouped_topk_cpu out-variant, 7 shapes x renormalize (:81-88)"),
    ("test_biased_grouped_topk.py", "biased_grouped_topk_cpu (:94-95)"),
    ("test_mxfp4.py", "mxfp4_scaled_mm_cpu + scale packing (:200-204)"),
    ("test_moe_offloading_cpu.py", "fused_experts with -1 / duplicate ids, fp8 and bf16 (:145)"),
]
BENCHES = [
    ("bench_moe.py", "bf16 / int8 / fp8 fused_experts at M = 4 and 3929, N 384, K 7168, E 256 (:144-145)"),
    ("bench_moe_offloading_cpu.py", "the same loop with -1 masked ids, M 64, N 256, K 4096, E 8 (:173)"),
]

VERDICT_RE = re.compile(r"(?:Comparing:|result:|test_weight_prepack:)\s+(True|False)\b|^\s*(?:tensor\()?(True|False)\)?\s*$")


def find_reference(arg):
    for cand in (arg, os.environ.get("SGLB_REFERENCE"), os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if cand and os.path.exists(os.path.join(cand, "test_moe.py")):
            return os.path.abspath(cand)
    return None


def run_script(ref, name, timeout):
    env = dict(os.environ)
    env["PYTHONPATH"] = DROPIN + os.pathsep + env.get("PYTHONPATH", "")
    t0 = time.time()
    try:
        p = subprocess.run([sys.executable, name], cwd=ref, env=env, capture_output=True, text=True, timeout=timeout)
        rc, out, err = p.returncode, p.stdout, p.stderr
    except subprocess.TimeoutExpired as e:
        rc, out, err = -9, (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or ""), "timeout"
    dt = time.time() - t0
    verdicts = []
    for line in out.splitlines():
        m = VERDICT_RE.search(line)
        if m:
            verdicts.append((m.group(1) or m.group(2)) == "True")
    return rc, out, err, verdicts, dt


def own_timing(shape_lines):
    """The build's own device-resident timing of the bench_moe.py shapes (bench_moe.py:144-145) next to the
    script's host-tensor loop: CUDA events over 20 calls after 3 warm-ups, weights rotated over 2 copies."""
    import torch
    sys.path.insert(0, ROOT)
    import bench
    import sgl_cpu_tests_b200 as K
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1111)
    E, topk, H, I = bench.SHAPES["refbench"]
    for M in (4, 3929):
        parts = []
        for fmt in ("bf16", "int8", "fp8"):
            ws = [bench.make_weights(fmt, E, H, I, dev, gen) for _ in range(2)]
            x, tw, ids = bench.make_batch(M, E, topk, H, dev, gen)

            def step(i):
                w = ws[i % 2]
                return K.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, bench.FMT_ID[fmt], w["s1"], w["s2"],
                                            w["block"], False)
            ms = bench.time_steps(step, 20, 3, 1) / 20
            parts.append(f"{fmt}: {ms:.3f} ms")
            del ws
            torch.cuda.empty_cache()
        shape_lines.append(f"### this build, device-resident inputs: M = {M}, N = {I}, K = {H}, E = {E}, TopK = {topk}: "
                           + "; ".join(parts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default=None)
    ap.add_argument("--out", default=None)
    ap.add_argument("--bench", action="store_true", help="also run bench_moe.py / bench_moe_offloading_cpu.py")
    ap.add_argument("--only", default=None, help="comma-separated script names")
    ap.add_argument("--timeout", type=int, default=1500)
    args = ap.parse_args()
    ref = find_reference(args.ref)
    if ref is None:
        print("reference scripts not found (pass --ref or set SGLB_REFERENCE)")
        return 2
    lines = [f"reference scripts from {ref}, run unmodified as `python3 <file>` with PYTHONPATH={os.path.relpath(DROPIN, ROOT)}"]
    failed = False
    todo = TESTS + (BENCHES if args.bench else [])
    if args.only:
        keep = set(args.only.split(","))
        todo = [t for t in todo if t[0] in keep]
    for name, what in todo:
        rc, out, err, verdicts, dt = run_script(ref, name, args.timeout)
        is_bench = name.startswith("bench_")
        n_true, n_false = sum(verdicts), len(verdicts) - sum(verdicts)
        ok = rc == 0 and n_false == 0 and (is_bench or n_true > 0)
        failed |= not ok
        lines.append(f"{'PASS' if ok else 'FAIL'}  {name:32s} rc={rc} verdicts: {n_true} True / {n_false} False  ({dt:.1f} s)  -- {what}")
        for l in out.splitlines():
            if VERDICT_RE.search(l) or l.startswith("### fused_experts"):
                lines.append("        | " + l.strip())
        if rc != 0:
            lines += ["        ! " + l for l in err.strip().splitlines()[-12:]]
    if args.bench:
        try:
            own_timing(lines)
        except Exception as e:  # the timing is a courtesy line, the verdicts above are the result
            lines.append(f"### this build's own timing failed: {e}")
    lines.append("RESULT: " + ("FAIL" if failed else "PASS"))
    text = "\n".join(lines)
    print(text)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            f.write(text + "\n")
    return 1 if failed else 0


if __name__ == "__main__":
    sys.exit(main())
