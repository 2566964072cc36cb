"""A/B pool-pipeline builds: usage pool_ab.py lib1 lib2 ... (final_scene 1080p x16, k_node 8 and 12)"""
import os, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
for lib in sys.argv[1:]:
    print(lib, flush=True)
    env = dict(os.environ, RTB200_LIB=str((ROOT / lib).resolve()))
    subprocess.run([sys.executable, str(ROOT / "tools" / "pool_sweep.py"), "final_scene", "1920", "1080", "16", "8", "12"], env=env, timeout=200)
