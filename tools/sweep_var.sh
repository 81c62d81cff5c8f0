# experiment helper: runs a command with an alternative build of the library (build/var/lib_X.so) in place of the product one
# usage: tools/sweep_var.sh X "command ..." [Y "command ..."] ...
mkdir -p gpurun_out; cp genome_weaver_align_b200/libgwalign.so /tmp/lib_orig.so
while [ $# -gt 0 ]; do
  v=$1; cmd=$2; shift 2
  cp build/var/lib_$v.so genome_weaver_align_b200/libgwalign.so
  echo "== $v: $cmd"; bash -c "$cmd"
done
cp /tmp/lib_orig.so genome_weaver_align_b200/libgwalign.so
