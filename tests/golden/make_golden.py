"""Writes tests/golden/haddock.json: the known-answer examples the reference carries in its
haddock comments (src/Data/Array/Accelerate/Numeric/LinearAlgebra.hs), transcribed by hand --
the reference is Haskell and cannot be imported or executed here, so these literals ARE the
fixture; this script only normalises them into one JSON file.  Expected values are the ones
printed in the reference's comments, not recomputed.
"""
import json
import os

CASES = [
    {"name": "dotu_real", "cite": "LinearAlgebra.hs:48-51", "op": "dotu", "dtype": "f64",
     "x": [1, 2, 3, 4], "y": [-2, 0, 1, 1], "expect": 5.0},
    {"name": "dotu_complex", "cite": "LinearAlgebra.hs:53-56", "op": "dotu", "dtype": "c64",
     "x": [[1, 1], [1, 0]], "y": [[1, 0], [1, -1]], "expect": [2.0, 0.0]},
    {"name": "outer", "cite": "LinearAlgebra.hs:65-71", "op": "outer", "dtype": "f64",
     "x": [1, 2, 3], "y": [5, 2, 3], "expect": [[5, 2, 3], [10, 4, 6], [15, 6, 9]]},
    {"name": "mXv", "cite": "LinearAlgebra.hs:86-95", "op": "mXv", "dtype": "f64",
     "m": [[1, 2, 3], [4, 5, 6]], "x": [10, 20, 30], "expect": [140.0, 320.0]},
    {"name": "vXm", "cite": "LinearAlgebra.hs:106-115", "op": "vXm", "dtype": "f64",
     "m": [[1, 2, 3], [4, 5, 6]], "x": [5, 10], "expect": [45.0, 60.0, 75.0]},
    {"name": "mXm", "cite": "LinearAlgebra.hs:129-149", "op": "mXm", "dtype": "f64",
     "a": [[1, 2, 3, 4, 5], [6, 7, 8, 9, 10], [11, 12, 13, 14, 15]],
     "b": [[1, 3], [0, 2], [-1, 5], [7, 7], [6, 0]],
     "expect": [[56, 50], [121, 135], [186, 220]]},
]

if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "haddock.json")
    with open(out, "w") as f:
        json.dump(CASES, f, indent=1)
    print("wrote", out)
