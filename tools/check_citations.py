#!/usr/bin/env python
"""Checks that every `file.ext:line[-line]` citation of the reference in our docs, headers and sources names a file that
exists under /root/reference and lines that exist in it (run in the build container only: the GPU box has no reference).
Usage: python tools/check_citations.py [paths...]; exits 1 if a citation is stale."""
import os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
PAT = re.compile(r"([A-Za-z0-9_\-./]+\.(?:cpp|hpp|h|c|cu|cuh|py|gadget|genic|sh|rst|example|rules)):(\d+)(?:\s*-\s*(\d+))?((?:,\s*\d+(?:\s*-\s*\d+)?)*)")
MORE = re.compile(r",\s*(\d+)(?:\s*-\s*(\d+))?")


def reference_files():
    out = {}
    for d, _, fs in os.walk(REF):
        if "/.git" in d:
            continue
        for f in fs:
            out.setdefault(f, []).append(os.path.join(d, f))
    return out


def main():
    paths = sys.argv[1:] or subprocess.check_output(["git", "ls-files"], cwd=ROOT, text=True).split()
    files = reference_files()
    nlines = {}
    bad = total = 0
    for p in paths:
        if not p.endswith((".md", ".h", ".cu", ".cuh", ".cpp", ".c", ".py")) or p.startswith(("SURVEY", "VERDICT", "ADVICE", "PAPERS", "SNIPPETS")):
            continue
        try:
            text = open(os.path.join(ROOT, p), errors="replace").read()
        except OSError:
            continue
        for m in PAT.finditer(text):
            name = m.group(1)
            ranges = [(int(m.group(2)), int(m.group(3) or m.group(2)))] + [(int(x), int(y or x)) for x, y in MORE.findall(m.group(4) or "")]
            base = os.path.basename(name)
            cands = [f for f in files.get(base, []) if f.endswith(name.lstrip("./"))] or files.get(base, [])
            if not cands:
                continue                                   # not a reference file (our own sources are cited too)
            for f in cands:
                if f not in nlines:
                    nlines[f] = sum(1 for _ in open(f, errors="replace"))
            for a, b in ranges:
                total += 1
                if not any(1 <= a <= b <= nlines[f] for f in cands):
                    bad += 1
                    print("%s: %s [%d-%d] -> no such lines (%s has %s)" % (p, m.group(0), a, b, base, ", ".join(str(nlines[f]) for f in cands)))
    print("%d reference citations checked, %d stale" % (total, bad))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
