"""Manual debugging aid for the GPU box: python tests/gpu_debug.py — runs the parity tests one by one and keeps going."""
import subprocess
import sys
r = subprocess.run([sys.executable, "-m", "pytest", "tests/test_gpu_parity.py", "-q", "-m", "gpu", "--tb=short", "-p", "no:cacheprovider"] + sys.argv[1:])
sys.exit(r.returncode)
