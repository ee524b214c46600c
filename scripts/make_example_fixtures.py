"""Cuts small fixtures out of the reference's own example reads (configs C1 and C2 of BASELINE.json) for the drop-in
test: the first 1200 pairs of example/reads/reads_{1,2}.fq and the first 500 reads of example/reads/longreads.fq that
are shorter than the checkpointing threshold (cminlen = 2000 rows, aligner_sw.cpp:514: the checkpointed DP is not on
the CUDA path).  Run where /root/reference is mounted; the outputs are committed under tests/golden/."""
import gzip
import os
import sys

REF = os.environ.get("BT2_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def records(path):
    with open(path) as fh:
        while True:
            rec = [fh.readline() for _ in range(4)]
            if not rec[0]:
                return
            yield rec


def main():
    for name in ("reads_1", "reads_2"):
        with gzip.open(os.path.join(OUT, "example_%s.fq.gz" % name), "wt") as out:
            for i, rec in enumerate(records(os.path.join(REF, "example", "reads", name + ".fq"))):
                if i >= 1200:
                    break
                out.writelines(rec)
    kept = 0
    with gzip.open(os.path.join(OUT, "example_longreads.fq.gz"), "wt") as out:
        for rec in records(os.path.join(REF, "example", "reads", "longreads.fq")):
            if len(rec[1].strip()) >= 2000:
                continue
            out.writelines(rec)
            kept += 1
            if kept >= 500:
                break
    # the whole of C2 (10000 pairs) and the whole of C1 except its 7 reads of >= 2000 rows: run through the batched server
    for name in ("reads_1", "reads_2"):
        with gzip.open(os.path.join(OUT, "example_%s_all.fq.gz" % name), "wt") as out:
            for rec in records(os.path.join(REF, "example", "reads", name + ".fq")):
                out.writelines(rec)
    with gzip.open(os.path.join(OUT, "example_longreads_all.fq.gz"), "wt") as out:
        for rec in records(os.path.join(REF, "example", "reads", "longreads.fq")):
            if len(rec[1].strip()) < 2000:
                out.writelines(rec)
    print("wrote fixtures to", OUT)


if __name__ == "__main__":
    sys.exit(main())
