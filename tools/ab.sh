#!/bin/bash
# time every build/variants/*.so on the cfg2 solve (tools/build_variant.sh makes them)
for so in build/variants/*.so; do
  echo "== $so: $(MVOLPS_LIB=$so python profiles/prof_solve.py --iters 100000 --reps 3 2>&1 | tail -1)"
done
