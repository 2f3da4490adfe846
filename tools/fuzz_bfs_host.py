"""Random configurations for the k_bfs CPU emulator (tests/bfs_host.bin, built by tests/test_bfs_core.py):
shapes, obstacle densities, slab counts, warp counts, map kinds and streamed-upload chunk sizes.

    python tools/fuzz_bfs_host.py [n] [seed]
"""
import os
import random
import subprocess
import sys

exe = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "bfs_host.bin")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
random.seed(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
failed = 0
for i in range(n):
    ranks = random.choice([1, 1, 1, 1, 2, 3, 4])
    X = random.choice([33, 40, 64, 70, 96, 130])
    Y = random.choice([8, 9, 16, 24, 30])
    Z = random.choice([8, 16, 24, 33, 40, 64])
    if ranks > 1:
        Z = max(Z, 8 * ranks + 1)
    args = [X, Y, Z, 1000 + i, random.choice([0, 5, 10, 20, 30, 40]), ranks, random.choice([1, 2, 3, 4]), random.choice([0, 0, 0, 1, 2, 3])]
    if ranks == 1 and random.random() < 0.5:
        args.append(random.choice([2, 3, 4, 5]))      # streamed upload, 2^k planes per chunk
    try:
        r = subprocess.run([exe] + [str(a) for a in args], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
        ok = r.returncode == 0 and r.stdout.startswith("OK") and " bad=0 " in r.stdout
        out = r.stdout[:300]
    except subprocess.TimeoutExpired:
        ok, out = False, "timeout"
    if not ok:
        failed += 1
        print("FAIL", args, out, flush=True)
print(f"ran {n}, failed {failed}")
sys.exit(1 if failed else 0)
