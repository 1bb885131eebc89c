# usage: tools/gpu/ncu_full.sh <out-name> <kernel-regex> <launch-skip> <cmd...>
# one `ncu --set full` capture of the first matching launch after <launch-skip> matches; report -> gpurun_out/<out-name>.ncu-rep
set -x
cd $GRAFT_REPO_ROOT
out=$1; kern=$2; skip=$3; shift 3
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$kern" -s $skip -c ${NCU_COUNT:-1} -f -o gpurun_out/$out "$@" > gpurun_out/$out.log 2>&1
tail -3 gpurun_out/$out.log
