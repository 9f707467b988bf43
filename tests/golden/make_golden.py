"""Generate tests/golden/ from the reference's own CPU translation units.

Run in the build container (needs /root/reference and `make -C oracle ref`):
    python tests/golden/make_golden.py
The reference ships no tests or golden vectors; these fixtures are
outputs of oracle/_ref/libgpusolve_ref.so, i.e. of src/cpu/*.cpp compiled unmodified.
golden.json   per case: level table, residual history (float.hex), compensated norms,
              sha256 of final fields, the centre value
fields.npz    full final fields of the small cases (bit-exact comparison off the box)
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from backends import OracleBackend, full_size_script, operator_script, V, F, R, NEWTONV  # noqa: E402
from cases import BIG_CASES, FULL_CASES, FULL_SAMPLE_STRIDE, OPERATOR_CASES, SOLVE_CASES, digest  # noqa: E402

KIND = "reference"


def solve_case(name, kw, arrays):
    b = OracleBackend(KIND, **kw)
    mode = kw["mode"]
    rec = {"kwargs": {k: (list(v) if isinstance(v, tuple) else v) for k, v in kw.items()},
           "num_levels": b.num_levels,
           "levels": [[list(b.level_info(l)[0]), b.level_info(l)[1].hex()] for l in range(b.num_levels)],
           "f0_digest": digest(b.get(0, F))}
    if mode == 2:
        norms, inner = b.newton_solve()
        rec["inner_cycles"] = inner
        sol = b.get(0, NEWTONV)
        rec["norm_exact_final"] = b.norm_exact(0, F).hex()
    else:
        norms = b.solve()
        sol = b.get(0, V)
        rec["norm_exact_final"] = b.norm_exact(0, R).hex()
    rec["norms"] = [x.hex() for x in norms]
    rec["norms_repr"] = [repr(x) for x in norms]
    rec["solution_digest"] = digest(sol)
    z, y, x = sol.shape
    rec["centre"] = float(sol[z // 2, y // 2, x // 2]).hex()
    rec["solution_absmax"] = float(np.max(np.abs(sol))).hex()
    if sol.size <= 40 ** 3:
        arrays[name + "/solution"] = sol
    b.close()
    return rec


def operator_case(name, kw, arrays):
    b = OracleBackend(KIND, maxiter=1, **kw)
    out = operator_script(b, kw["mode"])
    rec = {"kwargs": {k: (list(v) if isinstance(v, tuple) else v) for k, v in kw.items()}, "outputs": {}}
    for k, v in out.items():
        if isinstance(v, float):
            rec["outputs"][k] = v.hex()
        else:
            rec["outputs"][k] = digest(v)
            if v.size <= 24 ** 3:
                arrays[name + "/" + k] = v
    b.close()
    return rec


def full_case(name, kw, arrays):
    """BASELINE configs[1]-[3] at 513^3: norms, compensated norm of the last residual, digest and a strided sample"""
    b = OracleBackend(KIND, **kw)
    b.set_threads(os.cpu_count())
    norms, fid = full_size_script(b, kw["mode"], kw["maxiter"])
    sol = b.get(0, fid)
    rec = {"kwargs": {k: (list(v) if isinstance(v, tuple) else v) for k, v in kw.items()},
           "norms": [x.hex() for x in norms], "norms_repr": [repr(x) for x in norms],
           "norm_exact_final": b.norm_exact(0, R).hex(), "solution_digest": digest(sol),
           "solution_absmax": float(np.max(np.abs(sol))).hex(), "sample_stride": FULL_SAMPLE_STRIDE}
    st = FULL_SAMPLE_STRIDE
    arrays[name + "/sample"] = np.ascontiguousarray(sol[::st, ::st, ::st])
    b.close()
    return rec


def main():
    if "--missing" in sys.argv:  # add the records of cases that golden.json does not hold yet
        with open(os.path.join(HERE, "golden.json")) as f:
            gold = json.load(f)
        arrays = dict(np.load(os.path.join(HERE, "fields.npz")))
        for name, kw in {**SOLVE_CASES, **BIG_CASES}.items():
            if name not in gold["solve"]:
                gold["solve"][name] = solve_case(name, kw, arrays)
                print(name, gold["solve"][name]["norms_repr"][:3], flush=True)
        for name, kw in OPERATOR_CASES.items():
            if name not in gold["operators"]:
                gold["operators"][name] = operator_case(name, kw, arrays)
                print(name, flush=True)
        with open(os.path.join(HERE, "golden.json"), "w") as f:
            json.dump(gold, f, indent=1, sort_keys=True)
        np.savez_compressed(os.path.join(HERE, "fields.npz"), **arrays)
        return
    only_full = "--full-only" in sys.argv
    if only_full:  # add / refresh the 513^3 records without re-running everything else
        with open(os.path.join(HERE, "golden.json")) as f:
            gold = json.load(f)
        arrays = dict(np.load(os.path.join(HERE, "fields.npz")))
        gold["full"] = {}
        for name, kw in FULL_CASES.items():
            gold["full"][name] = full_case(name, kw, arrays)
            print(name, gold["full"][name]["norms_repr"], flush=True)
        with open(os.path.join(HERE, "golden.json"), "w") as f:
            json.dump(gold, f, indent=1, sort_keys=True)
        np.savez_compressed(os.path.join(HERE, "fields.npz"), **arrays)
        print("wrote", len(arrays), "arrays")
        return
    arrays = {}
    gold = {"generator": "tests/golden/make_golden.py", "source": "oracle/_ref/libgpusolve_ref.so (reference src/cpu/*.cpp)",
            "solve": {}, "operators": {}}
    for name, kw in {**SOLVE_CASES, **BIG_CASES}.items():
        gold["solve"][name] = solve_case(name, kw, arrays)
        print(name, gold["solve"][name]["norms_repr"][:3], flush=True)
    for name, kw in OPERATOR_CASES.items():
        gold["operators"][name] = operator_case(name, kw, arrays)
        print(name, flush=True)
    gold["full"] = {}
    for name, kw in FULL_CASES.items():
        gold["full"][name] = full_case(name, kw, arrays)
        print(name, gold["full"][name]["norms_repr"], flush=True)
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(gold, f, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "fields.npz"), **arrays)
    print("wrote", len(arrays), "arrays")


if __name__ == "__main__":
    main()
