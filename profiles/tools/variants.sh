# A/B of tuning variants of the CUDA library (built with `python -m pysdm_b200.build --out
# pysdm_b200/variants/<name>.so -D...`): kernel ms, ms per step, roofline fraction
for v in default "$@"; do
  if [ $v = default ]; then unset PYSDM_B200_SO; else export PYSDM_B200_SO=/root/repo/pysdm_b200/variants/$v.so; fi
  python bench.py --steps 100 --warmup 20 --no-cpu-baseline --no-e2e --no-configs 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', d['roofline']['kernel_ms'], d['ms_per_step'], d['roofline']['frac'])"
done
