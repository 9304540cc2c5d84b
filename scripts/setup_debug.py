"""Development aid: where does zkb_setup differ from the reference generator on a circuit?"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import zklaim_b200 as zk
from oracle import zkref as ref
from test_oracle_generator import randomness

ctx = zk.Context(0)
circ = ref.Circuit.sha_chain(1, seed=5)
d = circ.dims()
cs = dict(num_constraints=d["num_constraints"], num_variables=d["num_variables"], num_inputs=d["num_inputs"],
          csr_A=circ.csr(0), csr_B=circ.csr(1), csr_C=circ.csr(2))
toxic, g1, g2 = randomness(ref, 77)
circ.setup_fixed(toxic, g1, g2)
want = circ.export_key()
got = ctx.setup(cs, toxic, g1, g2)
print("dims", d, "swapped", got["swapped"])
for f in ("A_query", "B_index", "B_g1", "B_g2", "H_query", "L_query"):
    a, b = got[f], want[f]
    if a.shape != b.shape:
        print(f, "shape", a.shape, b.shape); continue
    if a.ndim == 1:
        print(f, "equal", np.array_equal(a, b)); continue
    bad = np.flatnonzero((a != b).any(axis=1))
    print(f, "rows", len(a), "mismatching", len(bad), bad[:10])
    for i in bad[:3]:
        h = a.shape[1] // 2
        print("   row", i, "var", int(got["B_index"][i]) if f.startswith("B_") else i, "x equal", np.array_equal(a[i, :h], b[i, :h]),
              "got zero", not a[i].any(), "want zero", not b[i].any())
        print("   got ", a[i]); print("   want", b[i])
