for w in 16777216 8388608 4194304; do
  python bench.py --wave-size $w --steps 4 --warmup 3 --no-cpu 2>/dev/null > gpurun_out/wave_$w.json
  python -c "import json; d=json.load(open('gpurun_out/wave_$w.json')); print($w, round(d['value'],1), round(d['ms_per_step'],1), d['waves_per_step'], round(d['e2e']['value'],1))"
done
