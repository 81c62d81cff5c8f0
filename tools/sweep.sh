# usage: tools/sweep.sh "opts1" "opts2" ...   (each a string of --opt name=value flags); prints one summary line per setting
mkdir -p gpurun_out
for o in "$@"; do
python bench.py --steps 2 --warmup 1 --no-cpu $o > gpurun_out/sw.log 2>gpurun_out/sw.err || tail -3 gpurun_out/sw.err
python tools/bench_line.py gpurun_out/sw.log "[$o]"
done
