#!/bin/bash
# usage: scratch/snaprun.sh <id> <outfile> <gpurun options...> -- <command run from the snapshot's root>
# Freezes a copy of the working tree under gpurun_snap/<id>/ and runs the GPU call from THAT copy, so that the tree can
# keep changing while the call waits for a box.  Retries while the pod answers busy.  The copy is removed afterwards.
id=$1; out=$2; shift 2
opts=()
while [ "$1" != "--" ]; do opts+=("$1"); shift; done
shift
cmd="$*"
root=$(cd "$(dirname "$0")/.." && pwd)
snap="$root/gpurun_snap/$id"
rm -rf "$snap"; mkdir -p "$snap"
( cd "$root" && tar cf - --exclude=./.git --exclude=./gpurun_out --exclude=./gpurun_snap --exclude='./oracle/_ref/obj_*' \
    --exclude=./of_dis_b200/build --exclude=./of_dis_b200/build_tol --exclude='*/__pycache__' --exclude=./.pytest_cache . ) | ( cd "$snap" && tar xf - )
for attempt in $(seq 1 60); do
  gpurun "${opts[@]}" -- "cd gpurun_snap/$id && mkdir -p gpurun_out && ( $cmd ); mkdir -p \$GRAFT_REPO_ROOT/gpurun_out && cp -r gpurun_out/. \$GRAFT_REPO_ROOT/gpurun_out/" > "$out" 2>&1
  rc=$?
  if grep -q "status=transient\|retry in a few minutes\|another call" "$out" || [ $rc -eq 3 ] || [ $rc -eq 2 ]; then sleep 60; continue; fi
  break
done
rm -rf "$snap"
echo "done rc=$rc attempt=$attempt" >> "$out"
