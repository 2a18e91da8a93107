#!/bin/bash
# Freeze a copy of the working tree under .stage/NAME (git-ignored, travels with gpurun) so that a GPU
# call which waits in the queue runs exactly this state while the tree keeps changing.
# usage: profiles/stage.sh NAME   ->  gpurun ... 'cd .stage/NAME && bash profiles/xyz.sh'
set -e
cd /root/repo
name=$1
rm -rf .stage/$name
mkdir -p .stage/$name
tar --exclude=./.git --exclude=./gpurun_out --exclude=./.stage --exclude='*.o' --exclude=__pycache__ --exclude=.pytest_cache \
    --exclude=./gsvc_b200/lib/variants/*.o -cf - . | tar -xf - -C .stage/$name
echo ".stage/$name ($(du -sh .stage/$name | cut -f1))"
