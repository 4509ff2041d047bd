timeout 300 python scripts/ln_probe.py 2>&1 | grep -v Warning | tail -16
