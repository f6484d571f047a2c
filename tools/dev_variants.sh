#!/bin/bash
# Dev tool: run a command once per variant library in _variants/ (copied over the in-tree library on the GPU box's snapshot).
for f in _variants/lib_*.so; do
  cp $f puflow_b200/_lib/libpuflow_b200.so
  echo "== $f"; "$@"
done
