"""One shape of the CTC sweep (tools/ctc_sweep.py): python tools/ctc_point.py T V L [B]   -- also the ncu target
(ncu --set full -k regex:ctc_ -c 4: prep, lattice, finish, grad of the first call)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ctc_sweep import run

T, V, L = (int(a) for a in sys.argv[1:4])
B = int(sys.argv[4]) if len(sys.argv) > 4 else 64
print(json.dumps(run(B, T, V, L, cpu_check=os.environ.get("CTC_CHECK") == "1")))
