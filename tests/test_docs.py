"""Doc lint: every entry point of the C ABI is accounted for in INTEGRATION.md / DESIGN.md, every
declaration in the header says what it replaces (or that nothing upstream corresponds), and the
files the docs point at exist."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = open(os.path.join(ROOT, "include", "b200validate.h")).read()
DOCS = {n: open(os.path.join(ROOT, n)).read() for n in ("DESIGN.md", "INTEGRATION.md", "README.md")}


def abi_functions():
    code = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
    return sorted(set(re.findall(r"\b(b200v_[a-z0-9_]+)\s*\(", code)))


def test_every_entry_point_is_documented():
    text = DOCS["INTEGRATION.md"] + DOCS["DESIGN.md"]
    missing = [f for f in abi_functions() if f not in text]
    assert missing == []


def test_header_states_provenance_of_each_group():
    # the words the header uses to tie a declaration to the reference (or to say there is no tie)
    assert HEADER.count("Replaces") + HEADER.count("replaces") >= 4
    assert "README.md:101-110" in HEADER and "README.md:152" in HEADER
    assert "no counterpart in the reference" in HEADER.replace("\n * ", " ")
    for row in ("a3", "a4", "a5", "a6", "a9", "a10"):
        assert "§8 %s" % row in HEADER or "§8(a) row" in HEADER, row


def test_docs_state_the_required_facts():
    d = DOCS["DESIGN.md"]
    assert "Parity is unpinned by the reference" in d
    assert "no CPU fallback" in d.replace("**", "")
    assert "Go is not installed" in d
    for section in ("## 1.", "## 2.", "## 3.", "## 4.", "## 5.", "## 6.", "## 7.", "## 8."):
        assert section in d
    i = DOCS["INTEGRATION.md"]
    assert 'import "C"' in i and "C.b200v_vectoradd" in i            # the cgo stub a maintainer adds


def test_paths_named_in_docs_exist():
    pat = re.compile(r"`((?:profiles|tools|tests|oracle|include|nvidia-gpu-operator-k8s-cluster_b200)/[A-Za-z0-9_./-]+)`")
    missing = []
    for name, text in DOCS.items():
        for path in pat.findall(text):
            if "{" in path or "*" in path or path.endswith("/"):
                continue
            if not os.path.exists(os.path.join(ROOT, path)):
                missing.append((name, path))
    # built artefacts are legitimately absent from a fresh checkout
    # (and oracle/_ref is named only to say that it cannot exist: the reference has no source)
    missing = [m for m in missing if "/bin/" not in m[1] and "/lib/" not in m[1] and m[1] != "oracle/_ref"]
    assert missing == []
