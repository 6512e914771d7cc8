#!/bin/bash
# Freeze HEAD (plus the built libraries) under gpurun_stage/<tag>/ so that a queued
# gpurun call runs a consistent tree while the working tree keeps changing:
#   scripts/stage.sh r02b && gpurun -- 'cd gpurun_stage/r02b && ln -sfn $GRAFT_REPO_ROOT/gpurun_out gpurun_out && bash scripts/...'
set -e
tag=$1
rm -rf gpurun_stage/$tag
mkdir -p gpurun_stage/$tag
git archive HEAD | tar -x -C gpurun_stage/$tag
cp mast3r_slam_quality_dualtsdf_b200/libb200tsdf.so gpurun_stage/$tag/mast3r_slam_quality_dualtsdf_b200/
[ -e oracle/liboracle.so ] && cp oracle/liboracle.so gpurun_stage/$tag/oracle/
if ls build/variants/*.so >/dev/null 2>&1; then mkdir -p gpurun_stage/$tag/build/variants && cp build/variants/*.so gpurun_stage/$tag/build/variants/; fi
[ -e MEASURED_PEAKS.json ] && cp MEASURED_PEAKS.json gpurun_stage/$tag/
# make the staged build look fresh (git archive sets mtimes to the commit time)
touch gpurun_stage/$tag/mast3r_slam_quality_dualtsdf_b200/libb200tsdf.so gpurun_stage/$tag/oracle/liboracle.so 2>/dev/null || true
echo staged gpurun_stage/$tag
