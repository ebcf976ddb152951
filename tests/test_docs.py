"""Documentation stays in step with the code: every environment switch the library reads is listed in DESIGN.md section 9,
and every profile file DESIGN.md / README.md / profiles/r02_SUMMARY.md cite exists."""
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _read(path):
    with open(os.path.join(ROOT, path), encoding="utf-8") as f:
        return f.read()


def test_environment_switches_are_documented():
    used = set()
    for pat in ("pointcloudprocessing_b200/csrc/*.cu", "pointcloudprocessing_b200/csrc/*.cuh", "pointcloudprocessing_b200/csrc/*.inl",
                "pointcloudprocessing_b200/*.py", "pointcloudprocessing_b200/*/*.py"):
        for path in glob.glob(os.path.join(ROOT, pat)):
            with open(path, encoding="utf-8") as f:
                src = f.read()
            used.update(re.findall(r'getenv\("(PCP_[A-Z0-9_]+)"\)', src))
            used.update(re.findall(r'environ(?:\.get)?\(?\[?"(PCP_[A-Z0-9_]+)"', src))
    design = _read("DESIGN.md")
    missing = sorted(v for v in used if v not in design)
    assert not missing, f"environment switches read by the code but not listed in DESIGN.md section 9: {missing}"


def test_cited_profile_files_exist():
    missing = []
    for doc in ("DESIGN.md", "README.md", "profiles/r02_SUMMARY.md"):
        text = _read(doc)
        for name in set(re.findall(r"`(?:profiles/)?(r0[12]_[A-Za-z0-9_.]+\.(?:json|csv|txt|log|md))`", text)):
            if not os.path.exists(os.path.join(ROOT, "profiles", name)):
                missing.append((doc, name))
    assert not missing, f"cited profile files that are not in profiles/: {sorted(missing)}"


def test_chain_kernel_hot_code_stays_small():
    """The production instantiation of the fused chain kernel is sized for the instruction cache (re-rolling its
    hot loops and moving trace hooks / time-out handlers out of them took it from 0.78 to 0.97 of the bf16 peak with the same
    pipeline).  Checked on the cross-compiled object: instruction count of chain2_kernel<false>, and no trace code in it."""
    import shutil
    import subprocess
    import pytest
    obj = os.path.join(ROOT, "pointcloudprocessing_b200", "csrc", "_obj", "chain_tc.o")
    if not os.path.exists(obj) or shutil.which("cuobjdump") is None:
        pytest.skip("chain_tc.o not built here / no cuobjdump")
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
    blocks = re.split(r"\n\s*Function : ", sass)
    prod = [b for b in blocks if b.startswith("_ZN3pcp13chain2_kernelILb0E")]
    assert len(prod) == 1, [b[:60] for b in blocks[1:]]
    body = prod[0]
    n_instr = len(re.findall(r"^\s*/\*[0-9a-f]{4,}\*/\s+\S", body, flags=re.M))
    assert 1500 < n_instr <= 3600, n_instr           # 5 248 before the re-roll, 3 312 after
    assert "SR_GLOBALTIMER" not in body              # trace_ev<false> leaves nothing behind
