"""Generate tests/golden/lamda_{test,co}.npz from the LAMDA text files that ship with the
reference (tests/data/test.txt, tests/data/co.txt). Run in the build container only
(/root/reference is not present on the GPU box); the .npz results are committed."""
import os, sys
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from magritte_b200 import lamda

REF = os.environ.get("MAGRITTE_REFERENCE", "/root/reference")
for name in ("test", "co"):
    ld = lamda.parse_lamda(os.path.join(REF, "tests", "data", f"{name}.txt"))
    lamda.save_linedata(os.path.join(HERE, f"lamda_{name}.npz"), ld)
    print(name, ld["nlev"], ld["nrad"], len(ld["colpar"]))
