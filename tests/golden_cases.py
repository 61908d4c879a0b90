import ast
import glob
import os
import numpy as np
from psascan_b200 import textgen

HERE = os.path.dirname(os.path.abspath(__file__))


def golden_cases():
    out = []
    for f in sorted(glob.glob(os.path.join(HERE, "golden", "*.npz"))):
        z = np.load(f, allow_pickle=False)
        kw = ast.literal_eval(str(z["kwargs"]))
        out.append((os.path.basename(f)[:-4], str(z["gen"]), kw, int(z["n"]), z["sa5"]))
    return out


def golden_text(gen, kw, n):
    return textgen.GENERATORS[gen](n, **kw)
