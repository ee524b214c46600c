"""CPU: the GPU-bound reference server has no CPU fallback.  Where no CUDA device is usable, the first read that reaches
the shim (bowtie2-server_b200/shim/) must end the run with bt2gpu's error -- not with alignments computed some other
way.  (On a GPU box this test does not apply and is skipped; tests/test_gpu_dropin.py covers that side.)"""
import os

import pytest

import dropin as D


def _cuda_usable():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:  # noqa: BLE001
        return False


@pytest.mark.skipif(_cuda_usable(), reason="a CUDA device is usable here")
def test_gpu_bound_server_fails_loudly_without_a_gpu(tmp_path):
    if not D.have_binaries():
        pytest.skip("oracle/_ref binaries missing: run __graft_entry__.build() where /root/reference is mounted")
    w = str(tmp_path)
    fa = os.path.join(w, "g.fa")
    with open(fa, "w") as fh:
        fh.write(">0\nTTGTTCGTTTGTTCGTACGATCGATCGGATCGATTAGCTAGCTAGGATCGAT\n")
    D.build_index(fa, os.path.join(w, "g"))
    fq = os.path.join(w, "g.fq")
    with open(fq, "w") as fh:
        fh.write("@r0\nTTGTTCGTTTGTTCGTACGATCGA\n+\n%s\n" % ("I" * 24))
    # the stock server aligns the read ...
    cpu, _ = D.run(D.CPU_SERVER, os.path.join(w, "g"), [], ["-U", fq], os.path.join(w, "cpu.sam"), threads=1, timeout=60)
    assert cpu and cpu[0].split("\t")[2] == "0"
    # ... the GPU-bound one must refuse
    with pytest.raises(RuntimeError) as ei:
        D.run(D.GPU_SERVER, os.path.join(w, "g"), [], ["-U", fq], os.path.join(w, "gpu.sam"), threads=1, timeout=60)
    assert "bt2gpu shim: bt2gpu_ctx_create" in str(ei.value)
    assert not os.path.exists(os.path.join(w, "gpu.sam")) or not [
        ln for ln in open(os.path.join(w, "gpu.sam")) if not ln.startswith("@") and ln.split("\t")[2] != "*"]
