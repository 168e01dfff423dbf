timeout 300 python tools/_os_prof.py 2>&1 | tail -12
