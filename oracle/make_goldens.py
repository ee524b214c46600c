#!/usr/bin/env python3
"""oracle/make_goldens.py -- TEST INFRASTRUCTURE.  Regenerates tests/golden/ by running the
UNMODIFIED reference built by oracle/Makefile (oracle/_ref/*) on the reference's own example
data (/root/reference/example).  Needs /root/reference, so it only runs in the build
container; the fixtures it writes are committed and travel to the GPU box.

  python oracle/make_goldens.py [--work /tmp/bt2_goldens]

DP goldens (tests/golden/dp_*.txt): problems the reference server itself logged with
--log-dp / --log-dp-opp (bt2_search.cpp:729-730) for configs C1/C2 in end-to-end and --local
mode, sampled, then replayed through the reference's SwAligner by oracle/ref_dp_dump.cpp
(P = problem, R = align() result + sorted candidates, A = each nextAlignment(), E = end state).
"""
import argparse
import os
import subprocess
import sys
import time
import zlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("BT2_REFERENCE", "/root/reference")
BIN = os.path.join(ROOT, "oracle", "_ref")
GOLD = os.path.join(ROOT, "tests", "golden")
MAX_CANDS_KEPT = 64


def run_server_client(work, name, server_args, client_args, port):
    logs = [os.path.join(work, "dp_%s.log" % name), os.path.join(work, "dpopp_%s.log" % name)]
    sam = os.path.join(work, "out_%s.sam" % name)
    err = open(os.path.join(work, "server_%s.err" % name), "w")
    srv = subprocess.Popen(
        [os.path.join(BIN, "bowtie2-align-server-s"), "--server-port", str(port), "-x",
         os.path.join(work, "lambda"), "-p", "1", "--seed", "0"] + server_args +
        ["--log-dp", logs[0], "--log-dp-opp", logs[1]],
        stdout=open(os.path.join(work, "server_%s.out" % name), "w"), stderr=err)
    try:
        for _ in range(300):
            txt = open(os.path.join(work, "server_%s.err" % name)).read()
            if "ready to process" in txt:
                break
            time.sleep(0.1)
        subprocess.check_call(
            [os.path.join(BIN, "bowtie2-align-l"), "--server-port", str(port), "-x",
             os.path.join(work, "lambda")] + client_args + ["-S", sam],
            stderr=open(os.path.join(work, "client_%s.err" % name), "w"))
    finally:
        srv.terminate()
        srv.wait()
    return logs, sam


def dump(args, out):
    with open(out, "w") as fh:
        subprocess.check_call([os.path.join(BIN, "ref_dp_dump")] + args, stdout=fh)


def sample_problems(dump_path, out_path, every, limit, maxlen=2000, want=None):
    """Keep every `every`-th P line (reads shorter than the checkpointing threshold cminlen=2000)."""
    kept = 0
    cur_p, cur_r = None, None
    idx = 0
    with open(dump_path) as fh, open(out_path, "w") as out:
        for ln in fh:
            if ln.startswith("P\t"):
                cur_p = ln
            elif ln.startswith("R\t") and cur_p is not None:
                t = cur_p.split("\t")
                r = ln.split("\t")
                take = (idx % every == 0)
                if want is not None and want(t, r):
                    take = True
                if take and len(t[1]) < maxlen and kept < limit:
                    out.write(cur_p)
                    kept += 1
                idx += 1
    return kept


def trim_cands(path):
    lines = []
    with open(path) as fh:
        for ln in fh:
            if ln.startswith("R\t"):
                t = ln.rstrip("\n").split("\t")
                if len(t) > 6 and t[6]:
                    t[6] = ";".join(t[6].split(";")[:MAX_CANDS_KEPT])
                ln = "\t".join(t) + "\n"
            lines.append(ln)
    with open(path, "w") as fh:
        fh.writelines(lines)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--work", default="/tmp/bt2_goldens")
    a = ap.parse_args()
    os.makedirs(a.work, exist_ok=True)
    os.makedirs(GOLD, exist_ok=True)
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-j8", "ref"],
                          stdout=subprocess.DEVNULL)
    subprocess.check_call([os.path.join(BIN, "bowtie2-build-s"),
                           os.path.join(REF, "example/reference/lambda_virus.fa"),
                           os.path.join(a.work, "lambda")], stdout=subprocess.DEVNULL)
    r1 = os.path.join(REF, "example/reads/reads_1.fq")
    r2 = os.path.join(REF, "example/reads/reads_2.fq")
    lr = os.path.join(REF, "example/reads/longreads.fq")
    runs = [
        ("c2", [], ["-1", r1, "-2", r2], []),
        ("c1", ["--sensitive"], ["-U", lr], []),
        ("c2loc", ["--local"], ["-1", r1, "-2", r2], ["--local"]),
        ("c1loc", ["--local"], ["-U", lr], ["--local"]),
    ]
    port = 15400
    for name, sargs, cargs, dargs in runs:
        port += 1
        logs, sam = run_server_client(a.work, name, sargs, cargs, port)
        for kind, log in (("ext", logs[0]), ("mate", logs[1])):
            if not os.path.exists(log) or os.path.getsize(log) == 0:
                continue
            full = os.path.join(a.work, "dump_%s_%s.txt" % (name, kind))
            dump(dargs + ["--log", log, "--index", os.path.join(a.work, "lambda")], full)
            probs = os.path.join(a.work, "probs_%s_%s.txt" % (name, kind))
            limit = 120 if name == "c1loc" else 400
            # always keep failures and 16-bit problems in the sample: they are the rare branches
            n = sample_problems(full, probs, every=23, limit=limit,
                                want=lambda t, r: r[1] == "0" and zlib.crc32(t[1].encode()) % 3 == 0)
            gold = os.path.join(GOLD, "dp_%s_%s.txt" % (name, kind))
            dump(dargs + ["--problems", probs], gold)
            trim_cands(gold)
            print("wrote", gold, n, "problems")
    return 0


if __name__ == "__main__":
    sys.exit(main())
