"""Transcribes the reference's PRNG-free known-answer tests into reference_kats.json.

Run in the build container (reads /root/reference, which does not exist on the GPU
box); the output is committed.  Only literal numbers are extracted from the test
sources -- no reference code is imported or copied.
"""
import ast, json, os, re, sys

REF = "/root/reference/tests"
out = {}

# tests/test_non_dominated_sort.py:8-28
out["nds_11x2"] = {
    "source": "tests/test_non_dominated_sort.py:8-28",
    "x": [[0.5, 4], [1, 2], [2, 1], [3, 1], [4, 1], [5, 0.5], [1, 4], [2, 3], [3, 2], [3, 3], [4, 4]],
    "rank": [0, 0, 0, 1, 2, 0, 1, 1, 2, 3, 4],
}
# tests/test_crowding_distance.py:7-21
out["crowd_5x2"] = {
    "source": "tests/test_crowding_distance.py:7-21",
    "x": [[0.5, 4], [1, 2.5], [2, 2], [3, 1], [4, 0.8]],
    "dist": ["inf", 1.5 / 3.5 + 2 / 3.2, 2 / 3.5 + 1.5 / 3.2, 2 / 3.5 + 1.2 / 3.2, "inf"],
}
# tests/test_crowding_distance.py:127-146
out["crowd_masked_9x2"] = {
    "source": "tests/test_crowding_distance.py:127-146",
    "x": [[-1, -1], [0.5, 4], [1, 2.5], [2, 2], [-2, -2], [3, 1], [4, 0.8], [-3, -3], [-3, -4]],
    "mask": [False, True, True, True, False, True, True, False, False],
    "dist": ["-inf", "inf", 1.5 / 3.5 + 2 / 3.2, 2 / 3.5 + 1.5 / 3.2, "-inf", 2 / 3.5 + 1.2 / 3.2, "inf", "-inf", "-inf"],
}
# tests/test_monitors.py:30-48 (front-0 extraction on int32 fitness, two rounds)
out["monitor_pf_int32"] = {
    "source": "tests/test_monitors.py:30-48",
    "fitness1": [[1, 2], [3, 1], [5, 6]],
    "pf1": [[1, 2], [3, 1]],
    "fitness2": [[0.5, 1.5], [7, 8], [0, 10]],
    "pf2": [[3, 1], [0.5, 1.5], [0, 10]],
}
# tests/test_crowding_distance.py:24-124: the 90 golden distances, parsed from the source text
src = open(os.path.join(REF, "test_crowding_distance.py")).read()
m = re.search(r"def test_crowding_distance2\(\):(.*?)def test_masked", src, re.S)
body = m.group(1)
arr = re.search(r"ground_truth = jnp\.array\(\s*\[(.*?)\]\s*\)", body, re.S).group(1)
vals = []
for tok in arr.split(","):
    tok = tok.strip()
    if not tok:
        continue
    vals.append("inf" if tok == "jnp.inf" else float(tok))
out["crowd_front0_128x8_prng314"] = {
    "source": "tests/test_crowding_distance.py:24-124",
    "input": "jax.random.normal(jax.random.PRNGKey(314), (128, 8)), rows with rank == 0",
    "rtol": 1e-5,
    "dist": vals,
}
# tests/test_lsmop.py:10-29: spot values of LSMOP1 on jax.random.uniform(PRNGKey(0), (100, 300), 0, upper)
out["lsmop1_100x300_prng0"] = {
    "source": "tests/test_lsmop.py:10-29",
    "input": "jax.random.uniform(PRNGKey(0), (100, 300), minval=0, maxval=[1, 1, 10 x 298]); LSMOP1(d=300, m=3)",
    "shape": [100, 3],
    "spot": [[1, 1, 12.5981], [1, 2, 21.2454]],
    "atol": 1e-4,
}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_kats.json")
json.dump(out, open(path, "w"), indent=1)
print("wrote", path, "with", len(vals), "golden distances")
