for s in 2 4 8 16; do
  LG_E2E_SLOTS=$s python bench.py --no-extra --no-pairs --steps 10 2>/dev/null > gpurun_out/e2e_$s.json
  python -c "import json; d=json.loads(open('gpurun_out/e2e_$s.json').read().strip().splitlines()[-1]); print('slots $s e2e', round(d['e2e']['value'],2), 'value', round(d['value'],1))"
done
