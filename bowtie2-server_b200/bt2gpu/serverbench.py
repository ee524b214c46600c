"""End-to-end harness around the reference's own server / client pair: the measurement (and parity) path of bench.py
and scripts/dropin_parity.py.

A *server* is one of
  oracle/_ref/bowtie2-align-server-s        the unmodified reference (the CPU arm),
  oracle/_ref/bowtie2-align-server-s-gpub   the same reference with SeedAligner / SwAligner / SwDriver::extend /
                                            GroupWalk2S bound to libbt2gpu.so through the BATCHED binding
                                            (bowtie2-server_b200/shim/batch_shim.cpp: worker fibers), our arm;
clients are the reference's own `bowtie2-align-l` (one process per read shard, as several users of one server).
Throughput is reads / wall time from the first client's start to the last client's exit: FASTQ parsing, the HTTP
framing, PCIe both ways, the reference's per-read driver, the sink and the SAM text are all inside.
"""
import os
import socket
import subprocess
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
BIN = os.path.join(ROOT, "oracle", "_ref")
CPU_SERVER = os.path.join(BIN, "bowtie2-align-server-s")                  # the reference's default build (-msse2)
CPU_SERVER_AVX2 = os.path.join(BIN, "avx2", "bowtie2-align-server-s")     # its SSE_AVX2=1 build (Makefile:79-83)
GPUB_SERVER = os.path.join(BIN, "bowtie2-align-server-s-gpub")
CLIENT = os.path.join(BIN, "bowtie2-align-l")


def have_binaries():
    return all(os.path.exists(p) for p in (CPU_SERVER, GPUB_SERVER, CLIENT))


def write_fastq(path, reads, quals, first_id=0, suffix=""):
    """reads/quals: [n, L] code / Phred+33 arrays -> FASTQ, names r<i><suffix>; vectorised (millions of reads)."""
    n, L = reads.shape
    c2a = np.frombuffer(b"ACGTN", dtype=np.uint8)
    names = [b"@r%d%s\n" % (first_id + i, suffix.encode()) for i in range(n)]
    body = np.empty((n, 2 * L + 4), dtype=np.uint8)
    body[:, :L] = c2a[reads]
    body[:, L] = 10; body[:, L + 1] = ord("+"); body[:, L + 2] = 10
    body[:, L + 3:2 * L + 3] = quals
    body[:, 2 * L + 3] = 10
    with open(path, "wb") as fh:
        rows = body.tobytes()
        w = 2 * L + 4
        fh.write(b"".join(names[i] + rows[i * w:(i + 1) * w] for i in range(n)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class Server:
    """One running server process."""

    def __init__(self, binary, base, workers, args=(), env=None, log=None, start_timeout=900):
        self.port = _free_port()
        self.log = log or os.path.join("/tmp", "bt2srv_%d.err" % self.port)
        e = dict(os.environ)
        e.update(env or {})
        self._err = open(self.log, "w")
        self.proc = subprocess.Popen([binary, "--server-port", str(self.port), "-x", base, "-p", str(workers), "--seed", "0"]
                                     + list(args), stdout=subprocess.DEVNULL, stderr=self._err, env=e)
        self.base = base
        t0 = time.time()
        while "ready to process" not in open(self.log).read():
            if self.proc.poll() is not None or time.time() - t0 > start_timeout:
                raise RuntimeError("server did not start: " + open(self.log).read()[-2000:])
            time.sleep(0.05)
        self.start_s = time.time() - t0

    def align(self, shards, out_prefix, timeout=3600):
        """shards: list of client argument lists (["-1", f1, "-2", f2] or ["-U", f]); one client process per shard,
        all started together.  Returns (wall seconds, [sam paths])."""
        procs, sams = [], []
        t0 = time.perf_counter()
        for k, cargs in enumerate(shards):
            sam = "%s_%d.sam" % (out_prefix, k)
            sams.append(sam)
            procs.append(subprocess.Popen([CLIENT, "--server-port", str(self.port), "-x", self.base] + list(cargs) + ["-S", sam],
                                          stdout=subprocess.DEVNULL, stderr=open(sam + ".err", "w")))
        deadline = time.time() + timeout          # one deadline for the whole pass, not one per client
        try:
            for p, sam in zip(procs, sams):
                rc = p.wait(timeout=max(0.1, deadline - time.time()))
                if rc != 0 or self.proc.poll() is not None:
                    raise RuntimeError("client exit %s; client stderr:\n%s\nserver stderr:\n%s"
                                       % (rc, open(sam + ".err").read()[-1500:], open(self.log).read()[-3000:]))
        finally:
            for p in procs:                       # a pass that failed or timed out leaves no client behind
                if p.poll() is None:
                    p.kill()
        return time.perf_counter() - t0, sams

    def rss_mb(self):
        try:
            for ln in open("/proc/%d/status" % self.proc.pid):
                if ln.startswith("VmRSS"):
                    return int(ln.split()[1]) / 1024.0
        except OSError:
            pass
        return None

    def close(self):
        if self.proc.poll() is None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=20)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        self._err.close()


def sam_records(paths):
    """All alignment lines of the given SAM files, sorted (record order depends on the threads, the records do not)."""
    out = []
    for p in paths:
        with open(p) as fh:
            out += [ln for ln in fh if not ln.startswith("@")]
    out.sort()
    return out


def count_differing(a, b):
    return sum(1 for x, y in zip(a, b) if x != y) + abs(len(a) - len(b))
