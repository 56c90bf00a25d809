TCVGG=1 timeout 300 python tools/tc_timeline.py 2>&1 | tail -12 | cut -c 1-220
