"""Diagnostic: total and gather-conv times out of scripts/eval_profile.py's last line (stdin)."""
import re
import sys
line = sys.stdin.read().strip().splitlines()[-1]
items = re.findall(r'(\S+_gather) (\d+\.\d+)x(\d+)', line)
tot = re.search(r'total (\d+\.\d+)', line)
print(tot.group(1), ' '.join(f'{n.replace("conv_", "")}={v}' for n, v, c in items))
