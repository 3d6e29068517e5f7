#!/usr/bin/env python3
"""Build oracle/_ref/libzvec_ref_<march>.so from the reference's OWN sources.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path (zvec_b200/, include/)
may load what this produces; only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs do.

Recipe: every *.cc under src/ailego except
ailego/buffer/buffer_manager.cc (needs Arrow), ailego/encoding/json/mod_json.c,
every *.cc of src/core/{framework,metric,quantizer,interface}, src/core/algorithm/
{flat,cluster,ivf} and src/core/utility except buffer_storage.cc, compiled where
they lie under /root/reference (never copied), plus oracle/ref_shim.cc (ours: a
C ABI over the reference's builder/searcher/converter classes) and a 3-method
link stub.  quantized_integer_metric.cc is built with -ffp-contract=off so the
int8 float epilogue is evaluated exactly as written (SURVEY §8c caveat 3).

Outputs go only to oracle/_ref/ (git-ignored, but shipped to the GPU box).
Two ISA levels are built: skylake-avx512 (parity oracle: fp32 accumulate for
fp16) and haswell (safe fallback when the host lacks AVX-512).
"""
import argparse
import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("ZVEC_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")


def sources():
    src = os.path.join(REF, "src")
    tus = []
    for root, _, files in os.walk(os.path.join(src, "ailego")):
        for f in files:
            p = os.path.join(root, f)
            if f.endswith(".cc") and not p.endswith("buffer/buffer_manager.cc"):
                tus.append(p)
    tus.append(os.path.join(src, "ailego/encoding/json/mod_json.c"))
    # core/interface = core_interface::Index / FlatIndex / IVFIndex / IndexFactory: the layer the DB calls
    # (vector_column_indexer.cc:195); it needs only the vendored header-only magic_enum
    for d in ("core/framework", "core/metric", "core/quantizer", "core/utility",
              "core/algorithm/flat", "core/algorithm/cluster", "core/algorithm/ivf",
              "core/interface", "core/interface/indexes"):
        for f in sorted(os.listdir(os.path.join(src, d))):
            if f.endswith(".cc") and f != "buffer_storage.cc":
                tus.append(os.path.join(src, d, f))
    return sorted(tus)


def compile_one(args):
    srcp, objp, march = args
    if os.path.exists(objp) and os.path.getmtime(objp) >= os.path.getmtime(srcp):
        return None
    os.makedirs(os.path.dirname(objp), exist_ok=True)
    src = os.path.join(REF, "src")
    inc = ["-I" + os.path.join(HERE, "stub"), "-I" + src, "-I" + os.path.join(src, "include"),
           "-I" + os.path.join(src, "core"), "-I" + os.path.join(src, "core/algorithm"),
           "-I" + os.path.join(REF, "thirdparty/magic_enum/magic_enum-0.9.7/include"),
           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "integration")]
    if srcp.endswith(".c"):
        cmd = ["gcc", "-O2", "-fPIC", "-march=" + march] + inc + ["-c", srcp, "-o", objp]
    else:
        cmd = ["g++", "-std=c++17", "-O2", "-fPIC", "-march=" + march, "-w"] + inc
        if srcp.endswith("quantized_integer_metric.cc"):
            cmd.append("-ffp-contract=off")
        cmd += ["-c", srcp, "-o", objp]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        return "FAILED %s\n%s" % (srcp, r.stderr[-3000:])
    return None


def build(march, jobs):
    objdir = os.path.join(OUT, "obj-" + march)
    tus = sources()
    own = [os.path.join(HERE, "stub/buffer_handle_stub.cc"), os.path.join(HERE, "ref_shim.cc")]
    work = []
    for s in tus:
        rel = os.path.relpath(s, os.path.join(REF, "src"))
        work.append((s, os.path.join(objdir, rel + ".o"), march))
    for s in own:
        work.append((s, os.path.join(objdir, "own", os.path.basename(s) + ".o"), march))
    with cf.ThreadPoolExecutor(jobs) as ex:
        errs = [e for e in ex.map(compile_one, work) if e]
    if errs:
        sys.stderr.write("\n".join(errs) + "\n")
        raise SystemExit(1)
    so = os.path.join(OUT, "libzvec_ref_%s.so" % march)
    objs = [w[1] for w in work]
    newest = max(os.path.getmtime(o) for o in objs)
    if not os.path.exists(so) or os.path.getmtime(so) < newest:
        subprocess.check_call(["g++", "-shared", "-o", so] + objs +
                              # --exclude-libs: keep the static libstdc++ private, so that loading this library
                              # RTLD_GLOBAL (for plugins) cannot interpose on torch's or anyone else's C++ runtime
                              ["-lpthread", "-ldl", "-lrt", "-static-libstdc++", "-static-libgcc",
                               "-Wl,--exclude-libs,ALL"])
    return so


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--march", nargs="*", default=["skylake-avx512", "haswell"])
    ap.add_argument("--jobs", type=int, default=os.cpu_count() or 4)
    a = ap.parse_args()
    if not os.path.isdir(os.path.join(REF, "src")):
        print("reference tree not present at %s: keeping prebuilt oracle/_ref" % REF)
        return
    os.makedirs(OUT, exist_ok=True)
    for m in a.march:
        print("built", build(m, a.jobs))


if __name__ == "__main__":
    main()
