"""Restates the reference's regression table (scripts/test/simple_tests.pl, a Perl data structure the GPU box cannot
run: no Perl modules, no /root/reference) as tests/golden/simple_cases.json: for every case that gives its reads
inline (`reads` or `mate1s`/`mate2s`), the reference sequences, the reads, the qualities, the option string and the
expected CIGARs where the table states them.  Run where /root/reference is mounted; the JSON is committed and
tests/test_gpu_dropin.py drives the stock and the GPU-bound reference servers with it."""
import json
import os
import re
import shlex
import sys

REF = os.environ.get("BT2_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "simple_cases.json")


def strip_comments(txt):
    out = []
    for ln in txt.split("\n"):
        q = False
        res = []
        i = 0
        while i < len(ln):
            c = ln[i]
            if c == "\\" and q:
                res.append(ln[i:i + 2]); i += 2; continue
            if c == '"':
                q = not q
            if c == "#" and not q:
                break
            res.append(c); i += 1
        out.append("".join(res))
    return "\n".join(out)


def blocks(txt):
    """Top-level { ... } blocks that contain a ref => key."""
    i = txt.index("my @cases")
    depth, start = 0, None
    q = False
    k = i
    while k < len(txt):
        c = txt[k]
        if c == "\\" and q:
            k += 2; continue
        if c == '"':
            q = not q
        elif not q:
            if c == "{":
                if depth == 0:
                    start = k
                depth += 1
            elif c == "}":
                depth -= 1
                if depth == 0 and start is not None:
                    yield txt[start:k + 1]
                    start = None
            elif c == ";" and depth == 0 and txt[k - 1] == ")":
                return
        k += 1


def perl_str(s):
    return s.replace('\\"', '"').replace("\\\\", "\\").replace("\\n", "\n").replace("\\t", "\t")


def strings(expr):
    """Double-quoted strings of a Perl list, `.`-concatenations joined."""
    toks = re.findall(r'"((?:[^"\\]|\\.)*)"|(\.)|(,)', expr)
    out, cur, cat = [], None, False
    for s, dot, comma in toks:
        if dot:
            cat = True
        elif comma:
            if cur is not None:
                out.append(cur)
            cur, cat = None, False
        else:
            cur = (cur + perl_str(s)) if (cat and cur is not None) else perl_str(s)
            cat = False
    if cur is not None:
        out.append(cur)
    return out


def listval(block, key):
    m = re.search(r"\b%s\s*=>\s*\[" % key, block)
    if not m:
        return None
    k, depth = m.end(), 1
    while depth:
        depth += {"[": 1, "]": -1}.get(block[k], 0)
        k += 1
    return strings(block[m.end():k - 1])


def scalar(block, key):
    m = re.search(r'\b%s\s*=>\s*((?:"(?:[^"\\]|\\.)*"\s*\.?\s*)+)' % key, block)
    if not m:
        m = re.search(r"\b%s\s*=>\s*(\d+)" % key, block)
        return m.group(1) if m else None
    return "".join(strings(m.group(1)))


def runs_on_stock_reference(c):
    """Keep the cases the stock server/client pair (oracle/_ref) accepts with these options and that do not ask for
    side files.  (Seeds with mismatches -- -N 1, SEED=1/2, --multiseed=1 -- are on the CUDA path since round 2.)"""
    import tempfile
    sys.path.insert(0, os.path.join(os.path.dirname(OUT), ".."))
    import dropin as D
    a = " ".join(c["args"])
    if re.search(r"--un|--al|-p ", a):
        return False
    work = tempfile.mkdtemp(prefix="simple_")
    try:
        base, sargs, cargs = D.write_case_inputs(c, work, "c")
        lines, _ = D.run(D.CPU_SERVER, base, sargs, cargs, os.path.join(work, "o.sam"), threads=1, timeout=60)
        return len(lines) > 0
    except Exception as e:  # noqa: BLE001 - any failure of the stock run disqualifies the case
        print("dropped %r: %s" % (c["name"], str(e).splitlines()[0][:100]))
        return False


def main():
    txt = strip_comments(open(os.path.join(REF, "scripts", "test", "simple_tests.pl")).read())
    cases, skipped = [], 0
    for b in blocks(txt):
        if not re.search(r"\bref\s*=>", b):
            continue
        ref, reads = listval(b, "ref"), listval(b, "reads")
        m1, m2 = listval(b, "mate1s"), listval(b, "mate2s")
        # reads given as the CONTENT of a read file (the table's parser cases): kept as content + format
        files = None
        for fmt, flag in (("fastq", "-q"), ("tabbed", "--tab5"), ("fasta", "-f"), ("qseq", "--qseq"), ("raw", "-r")):
            one, f1, f2 = scalar(b, fmt), scalar(b, fmt + "1"), scalar(b, fmt + "2")
            if one is not None and not one.isdigit():
                files = {"flag": flag, "file": one}
            elif f1 is not None and f2 is not None:
                files = {"flag": flag, "file1": f1, "file2": f2}
        if not ref or (not reads and not (m1 and m2) and not files) or re.search(r"\b(should_abort|idx_map|color)\s*=>", b):
            skipped += 1
            continue
        args = (scalar(b, "args") or "") + " " + (scalar(b, "report") or "")
        try:
            argv = [a.replace("\\;", ";") for a in shlex.split(args)]
        except ValueError:
            skipped += 1
            continue
        c = {"name": scalar(b, "name") or "case%d" % len(cases), "ref": ref, "args": argv}
        if reads:
            c["reads"] = reads
            q = listval(b, "quals")
            if q:
                c["quals"] = q
        elif files and not (m1 and m2):
            c["files"] = files
        else:
            c["mate1s"], c["mate2s"] = m1, m2
            f1, f2 = scalar(b, "mate1fw"), scalar(b, "mate2fw")
            c["orient"] = {("1", "0"): "--fr", ("1", "1"): "--ff", ("0", "1"): "--rf", ("0", "0"): "--ff"}[(f1 or "1", f2 or "0")]
        cig = listval(b, "cigar")
        if cig:
            c["cigar"] = cig
        for k in ("norc", "nofw"):
            if scalar(b, k) == "1":
                c[k] = 1
        cases.append(c)
    cases = [c for c in cases if runs_on_stock_reference(c)]
    json.dump(cases, open(OUT, "w"), indent=0)
    print("%d cases restated, %d skipped -> %s" % (len(cases), skipped, OUT))


if __name__ == "__main__":
    sys.exit(main())
