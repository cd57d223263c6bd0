"""Extracts the reference's Rust-generated field golden vectors into a compact fixture.

Source : /root/reference/test-vectors/{m31,cm31,qm31,securecolumn}-test-vectors.json
         (replayed by the reference at test-equivalence/fields/*.test.ts)
Output : tests/golden/field_vectors.json  (only operation / inputs / output; intermediates dropped)
Run in the build container only (the GPU box has no /root/reference); the output is committed.
"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/test-vectors"
out = {}
for f in ["m31", "cm31", "qm31", "securecolumn"]:
    d = json.load(open(f"{SRC}/{f}-test-vectors.json"))
    vs = []
    for v in d["test_vectors"]:
        inp = {k: x for k, x in v["inputs"].items() if k != "test_case"}
        vs.append({"op": v["operation"], "in": inp, "out": v["output"]})
    out[f] = vs
dst = os.path.join(ROOT, "tests", "golden", "field_vectors.json")
json.dump(out, open(dst, "w"), separators=(",", ":"))
print(dst, {k: len(v) for k, v in out.items()}, os.path.getsize(dst), "bytes")
