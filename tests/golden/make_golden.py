"""Regenerates the committed golden fixtures from the reference tree (run in the authoring
container, where /root/reference exists; the GPU box only ever sees the .npz files).

  signapprox.npz : tools/figures/signApprox.csv columns `input`, `combined` (the checked-in output of
                   OpenFHEWrapper::chebyshevCompare(delta=0.44, depth=10)), every 8th row plus every
                   row with input in [0.40, 0.50] (the decision region).
  dat_2_10.npz   : test/2_10.dat (n, query, database) as int8, the reference's shipped smoke dataset
                   (run_artifact.sh:7-27); expected result membership = true, index = [0].
"""
import csv
import os

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))

rows = list(csv.reader(open(os.path.join(REF, "tools/figures/signApprox.csv"))))[1:]
x = np.array([float(r[0]) for r in rows])
comb = np.array([float(r[1]) for r in rows])
keep = np.zeros(len(x), dtype=bool)
keep[::8] = True
keep |= (x >= 0.40) & (x <= 0.50)
np.savez_compressed(os.path.join(HERE, "signapprox.npz"), input=x[keep], combined=comb[keep])

tok = open(os.path.join(REF, "test/2_10.dat")).read().split()
n = int(tok[0])
vals = np.array(tok[1:], dtype=np.int64)
assert vals.min() >= -128 and vals.max() <= 127
np.savez_compressed(os.path.join(HERE, "dat_2_10.npz"), n=n, query=vals[:512].astype(np.int8),
                    db=vals[512:512 + n * 512].reshape(n, 512).astype(np.int8))
print("signapprox rows:", int(keep.sum()), " 2_10 n:", n)
