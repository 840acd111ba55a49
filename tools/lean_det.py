import os, sys
from pathlib import Path
import numpy as np
sys.path.insert(0, str(Path(__file__).resolve().parent))
os.environ["LEAN_ONLY_K23"] = "1"
import importlib
ld = importlib.import_module("lean_diff")
for name, env in (("lean", {}), ("general", {"B200CFD_K23_LEAN": "0"})):
    a = ld.run(name + "1", env); b = ld.run(name + "2", env)
    print(name, "run-to-run differing cells:", {k: int((a[k] != b[k]).sum()) for k in ("div", "phi", "u", "v")})
