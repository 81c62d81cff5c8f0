for o in "--opt wave_reads=1048576" "--opt wave_reads=1048576 --opt tail_steps=16" "--opt wave_reads=1048576 --opt tail_steps=16 --opt search_blocks=592" "--opt wave_reads=1048576 --opt tail_steps=8 --opt search_blocks=592" "--opt wave_reads=1048576 --opt tail_steps=32 --opt search_blocks=296"; do
python bench.py --steps 1 --warmup 1 --no-cpu $o > gpurun_out/sw.log 2>gpurun_out/sw.err || tail -3 gpurun_out/sw.err
python -c "
import json; d=json.loads(open('gpurun_out/sw.log').read().strip().splitlines()[-1]); print('$o', round(d['value']), round(d['ms_per_step'],1), {k:round(v['ms'],2) for k,v in d['roofline']['families'].items()}, d['counters_per_step']['rounds'])"
done
