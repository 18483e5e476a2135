"""Runs bench.py with the given extra arguments and prints ms_per_step / value only (developer helper)."""
import json
import subprocess
import sys

out = subprocess.run([sys.executable, "bench.py", "--no-sweep", "--no-cpu-baseline"] + sys.argv[1:], capture_output=True, text=True)
for line in out.stdout.splitlines():
    if line.startswith("{"):
        j = json.loads(line)
        print(f"ms_per_step {j['ms_per_step']:.4f}  value {j['value']:.4g}  tokenize_ms {j.get('tokenize', {}).get('ms_per_step', float('nan')):.4f}"
              f"  search_us {j.get('roofline', {}).get('us_per_launch', float('nan')):.2f}")
        break
else:
    print(out.stdout[-2000:], out.stderr[-2000:])
