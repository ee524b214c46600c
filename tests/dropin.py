"""Drop-in harness (test infrastructure): runs the unmodified reference server/client pair (oracle/_ref, built from
/root/reference by oracle/Makefile) and the same reference linked against libbt2gpu.so through the shim in
bowtie2-server_b200/shim/ (oracle/_ref/bowtie2-align-server-s-gpu) on the same reads, and returns both SAM outputs."""
import os
import socket
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, os.path.join(ROOT, "bowtie2-server_b200"))

CPU_SERVER = os.path.join(BIN, "bowtie2-align-server-s")
GPU_SERVER = os.path.join(BIN, "bowtie2-align-server-s-gpu")      # per-call binding (shim/gpu_shim.cpp)
GPUB_SERVER = os.path.join(BIN, "bowtie2-align-server-s-gpub")    # batched binding: worker fibers (shim/batch_shim.cpp)
EMUL_SERVER = os.path.join(BIN, "bowtie2-align-server-s-emul")    # test double: batched binding over the CPU oracle (oracle/emul_abi.c)
CLIENT = os.path.join(BIN, "bowtie2-align-l")


def have_binaries():
    return all(os.path.exists(p) for p in (CPU_SERVER, GPU_SERVER, GPUB_SERVER, CLIENT, os.path.join(BIN, "bowtie2-build-s")))


def read_fasta(path):
    names, seqs, cur = [], [], []
    for ln in open(path):
        if ln.startswith(">"):
            if cur:
                seqs.append("".join(cur)); cur = []
            names.append(ln[1:].strip())
        else:
            cur.append(ln.strip())
    seqs.append("".join(cur))
    return names, seqs


def build_index(fasta, base, threads=1):
    subprocess.check_call([os.path.join(BIN, "bowtie2-build-s"), "-q", "--threads", str(threads), fasta, base],
                          stdout=subprocess.DEVNULL)


def write_fastq(path, reads, quals, prefix, lens=None):
    """reads/quals: [n, L] code / phred+33 arrays; lens: optional per-read length (ragged)."""
    with open(path, "w") as fh:
        for i in range(len(reads)):
            n = reads.shape[1] if lens is None else int(lens[i])
            fh.write("@%s%d\n%s\n+\n%s\n" % (prefix, i, "".join("ACGTN"[c] for c in reads[i, :n]),
                                               bytes(quals[i, :n]).decode()))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run(server, base, server_args, client_args, sam, threads=2, timeout=600, env=None):
    """One server + one client run; returns (sorted alignment lines, server stderr)."""
    port = _free_port()
    errp, outp = sam + ".server.err", sam + ".server.out"
    e = dict(os.environ)
    e.update(env or {})
    with open(errp, "w") as err, open(outp, "w") as out:
        srv = subprocess.Popen([server, "--server-port", str(port), "-x", base, "-p", str(threads), "--seed", "0"]
                               + server_args, stdout=out, stderr=err, env=e)
        try:
            t0 = time.time()
            while "ready to process" not in open(errp).read():
                if srv.poll() is not None or time.time() - t0 > 120:
                    raise RuntimeError("server did not start: " + open(errp).read()[-2000:])
                time.sleep(0.05)
            rc = subprocess.call([CLIENT, "--server-port", str(port), "-x", base] + client_args + ["-S", sam],
                                 stderr=open(sam + ".client.err", "w"), timeout=timeout)
            if rc != 0:
                raise RuntimeError("client exit %d; client stderr:\n%s\nserver stderr:\n%s"
                                   % (rc, open(sam + ".client.err").read()[-1500:], open(errp).read()[-1500:]))
        finally:
            srv.terminate()
            try:
                srv.wait(timeout=30)
            except subprocess.TimeoutExpired:
                srv.kill()
    lines = sorted(ln for ln in open(sam) if not ln.startswith("@"))
    return lines, open(errp).read()


def genome_codes(fasta="lambda.fa"):
    """All sequences of a golden FASTA as code arrays (0..3, 4 = N)."""
    _, seqs = read_fasta(os.path.join(GOLDEN, fasta))
    lut = np.full(256, 4, np.uint8)
    for i, c in enumerate("ACGT"):
        lut[ord(c)] = i
        lut[ord(c.lower())] = i
    return [lut[np.frombuffer(x.encode(), np.uint8)] for x in seqs]


def lambda_genome():
    return genome_codes("lambda.fa")[0]


def write_case_inputs(case, work, tag):
    """Index + read files of one restated regression case (tests/golden/simple_cases.json).  Like the reference's
    runner, unpaired reads are given forward and reverse-complemented (unless the case says nofw / norc).
    Returns (index base, server args, client args)."""
    base = os.path.join(work, tag)
    with open(base + ".fa", "w") as fh:
        for i, r in enumerate(case["ref"]):
            fh.write(">%d\n%s\n" % (i, r))
    build_index(base + ".fa", base)
    comp = str.maketrans("ACGTNacgtn", "TGCANtgcan")
    args = list(case["args"])
    if "reads" in case:
        fq = base + ".fq"
        with open(fq, "w") as fh:
            for i, r in enumerate(case["reads"]):
                q = case["quals"][i] if "quals" in case and i < len(case["quals"]) else "I" * len(r)
                if not case.get("nofw"):
                    fh.write("@r%d\n%s\n+\n%s\n" % (i, r, q))
                if not case.get("norc"):
                    fh.write("@rc%d\n%s\n+\n%s\n" % (i, r.translate(comp)[::-1], q[::-1]))
        return base, args, ["-U", fq]
    if "files" in case:          # the table's read-file cases: content and format flag as the Perl runner passes them
        fl = case["files"]
        if "file" in fl:
            f = base + ".reads"
            open(f, "w").write(fl["file"])
            return base, args, [fl["flag"], f] if fl["flag"] == "--tab5" else [fl["flag"], "-U", f]
        f1, f2 = base + "_1.reads", base + "_2.reads"
        open(f1, "w").write(fl["file1"]); open(f2, "w").write(fl["file2"])
        return base, args, [fl["flag"], "-1", f1, "-2", f2]
    f1, f2 = base + "_1.fq", base + "_2.fq"
    for f, ms in ((f1, case["mate1s"]), (f2, case["mate2s"])):
        with open(f, "w") as fh:
            for i, r in enumerate(ms):
                fh.write("@p%d\n%s\n+\n%s\n" % (i, r, "I" * len(r)))
    return base, args + [case["orient"]], ["-1", f1, "-2", f2]
