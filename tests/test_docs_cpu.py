"""Docs stay true: files and tests that the documentation names exist."""
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
PKG = "csed_514_project_distributed_training_using_pytorch_b200"


def _backticked(text):
    return re.findall(r"`([^`\n]+)`", text)


def test_files_named_in_docs_exist():
    missing = []
    for doc in ("docs/COVERAGE.md", "profiles/README.md", "scripts/README.md", "DESIGN.md",
                "INTEGRATION.md", "README.md", "docs/ROADMAP.md"):
        base = (ROOT / doc).parent
        for tok in _backticked((ROOT / doc).read_text()):
            tok = tok.split("::")[0].strip()
            if not re.fullmatch(r"[\w./{},*-]+\.(py|cu|cuh|cpp|h|md|json|csv|sh)", tok):
                continue
            if any(c in tok for c in "{*"):
                continue                                   # patterns like final_n{1,2}_*.json
            cands = [ROOT / tok, base / tok, ROOT / tok.replace("pkg/", PKG + "/"),
                     ROOT / PKG / tok, ROOT / "scripts" / tok, ROOT / "profiles" / tok,
                     ROOT / "csrc" / tok, ROOT / "tests" / tok]
            if tok.startswith(("src/", "torch/", "baseline/", "lib/")) or tok in ("MEASURED_PEAKS.json",
                                                                               "BASELINE.json"):
                continue                                   # reference / torch / generated paths
            if not any(c.exists() for c in cands):
                missing.append(f"{doc}: {tok}")
    assert not missing, missing


def test_tests_named_in_coverage_exist():
    names = set()
    for f in (ROOT / "tests").glob("test_*.py"):
        names.update(re.findall(r"^def (test_\w+)", f.read_text(), flags=re.M))
    cited = set()
    for tok in _backticked((ROOT / "docs" / "COVERAGE.md").read_text()):
        for t in re.findall(r"test_\w+", tok.split("::")[-1] if "::" in tok else tok):
            if not t.endswith("_") and not (ROOT / "tests" / (t + ".py")).exists():
                cited.add(t)                                 # skip prefixes (test_x_*) and file names
    unknown = sorted(t for t in cited if t not in names and not any(n.startswith(t) for n in names))
    assert not unknown, unknown
