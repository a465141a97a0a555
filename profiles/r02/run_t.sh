# r02 call T2: resize tile width x CTAs per SM
for tl in 7 6 5; do for cps in 6 8; do echo "== tile 2^$tl CTAs/SM $cps"; ICB_RS_TL=$tl ICB_RS_CPS=$cps timeout 300 python profiles/r02/resize_probe.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('  ', d['src'], d.get('alpha', 'semi' if d['src'][2] == 4 else '-'), d['resize_us'], d['frac_of_measured_peak'])"; done; done
