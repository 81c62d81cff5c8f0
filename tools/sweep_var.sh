# experiment helper: runs bench.py with alternative builds of the library (build/var/lib_X.so) and per-build options
mkdir -p gpurun_out; cp genome_weaver_align_b200/libgwalign.so /tmp/lib_orig.so
while [ $# -gt 0 ]; do
  v=$1; o=$2; shift 2
  cp build/var/lib_$v.so genome_weaver_align_b200/libgwalign.so
  python bench.py --steps 2 --warmup 1 --no-cpu $o > gpurun_out/sw.log 2>gpurun_out/sw.err || tail -3 gpurun_out/sw.err
  python tools/bench_line.py gpurun_out/sw.log "[$v $o]"
done
cp /tmp/lib_orig.so genome_weaver_align_b200/libgwalign.so
