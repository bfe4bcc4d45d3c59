python -m pytest tests/test_slices.py tests/test_host_cpp.py -m gpu -x -q 2>&1 | tail -2
python tools/bench_configs.py 2>&1 | grep slices | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['config'][:5], d['emit_ms'], d['emit_frac_of_hbm_peak'])"
