import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import zecale_b200 as z
from oracle import ref as R
ref = R.Ref()
ctx = z.Context(0)
n = int(sys.argv[2])
scen = sys.argv[1]
P = ref.bases_mul(ref.fr_rng(10, n), 1)
Q = ref.bases_mul(ref.fr_rng(20, n), 1)
inf = np.zeros(36, dtype=np.uint64); inf[12:24] = P[0][24:36]
op = 0
if scen == "dbl": Q[0] = P[0]
if scen == "neg": Q[1] = ref.group_op(4, P[1:2], None, 1)[0]
if scen == "infa": P[2] = inf
if scen == "infb": Q[3] = inf
if scen == "mdbl": Q[0] = P[0]; op = 1
norm = lambda x: ref.group_op(3, x, None, 1)
try:
    got = ctx.test_group(op, P, Q)
    exp = norm(ref.group_op(op, P, Q, 1))
    print(scen, n, "equal:", np.array_equal(got, exp), flush=True)
except Exception as e:
    print(scen, n, "EXC", str(e)[-60:], flush=True)
