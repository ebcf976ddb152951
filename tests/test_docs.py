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
