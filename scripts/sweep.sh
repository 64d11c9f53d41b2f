python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for L in 1 2 4 8; do echo "LANES=$L"; SPGX_VBD_LANES=$L python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()); continue
    print('ms/step',round(d['ms_per_step'],3),'value %.3e'%d['value'],'frac',round(d['roofline']['frac'],3),'e2e %.3e'%d['e2e']['value'])
"; done
echo NOELL; SPGX_NO_ELL=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()); continue
    print('ms/step',round(d['ms_per_step'],3),'value %.3e'%d['value'],'frac',round(d['roofline']['frac'],3))
"
