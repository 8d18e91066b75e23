"""Every evidence file the documents cite under profiles/ exists, is not empty and,
if it is a .json, parses (two cited bench files were zero bytes)."""
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DOCS = ["profiles/README.md", "DESIGN.md", "README.md", "INTEGRATION.md"]


def cited():
    names = set()
    for d in DOCS:
        path = os.path.join(ROOT, d)
        if not os.path.exists(path):
            continue
        text = open(path).read()
        names |= set(re.findall(r"`(?:profiles/)?(r\d+_[A-Za-z0-9_.\-]+\.(?:json|csv|txt|ncu-rep|log))`", text))
    return sorted(names)


def test_cited_evidence_files_exist_and_are_not_empty():
    names = cited()
    assert len(names) > 10
    missing = [n for n in names if not os.path.exists(os.path.join(ROOT, "profiles", n))]
    assert not missing, f"cited but absent: {missing}"
    empty = [n for n in names if os.path.getsize(os.path.join(ROOT, "profiles", n)) == 0]
    assert not empty, f"cited but empty: {empty}"
    for n in names:
        if n.endswith(".json"):
            with open(os.path.join(ROOT, "profiles", n)) as f:
                lines = [l for l in f.read().splitlines() if l.strip().startswith("{")]
            assert lines, f"{n}: no JSON line"
            for l in lines:
                json.loads(l)


def test_no_empty_file_under_profiles():
    for n in os.listdir(os.path.join(ROOT, "profiles")):
        assert os.path.getsize(os.path.join(ROOT, "profiles", n)) > 0, n
