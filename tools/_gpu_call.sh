timeout 300 python tools/_tmp_occ.py 2>&1 | tail -2
