#!/bin/bash
# quick phase timing of one workload for each value of an env knob:  bash scripts/phase_times.sh VKM_K1 0 1 2
knob=$1; shift
for v in "$@"; do echo "== $knob=$v"; env $knob=$v python scripts/profile_import.py --imports 6 | tail -1; done
