for g in 2 4 6 8 13 26; do
  echo "groups=$g"; EMB_GROUPS=$g python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());e=d['e2e'];print(round(e['ms_per_step'],3), e['last_call_ms'], {k:round(v['ms_per_step'],3) for k,v in e['other_paths'].items()})"
done
