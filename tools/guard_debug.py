"""Debug helper: run the guard test's driver under B2_GUARD=1 with progress prints."""
import os, sys, re
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import importlib.util
spec = importlib.util.spec_from_file_location("tg", os.path.join(ROOT, "tests", "test_gpu_guard.py"))
src = open(os.path.join(ROOT, "tests", "test_gpu_guard.py")).read()
drv = src.split('DRIVER = r"""')[1].split('"""')[0]
drv = drv.replace("    E.update(obs, a, lp, rw, fl)\n", "    print('update', B, lam, flush=True)\n    try:\n        E.update(obs, a, lp, rw, fl)\n    except Exception as ex:\n        import numpy as np\n        print('FAILED', ex, 'losses', E.last_update_losses(2), 'a finite', np.isfinite(a).all(), np.isfinite(lp).all(), 'ret', np.isfinite(E.traj_download(1, fields=('returns',))['returns']).all() if False else '')\n        raise\n")
sys.argv = ["x", ROOT]
exec(compile(drv, "driver", "exec"))
