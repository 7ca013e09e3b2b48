#!/bin/bash
# A/B timing of kernel variants on the bench workload (development tool): ab.sh tag1 tag2 ...
cd "$(dirname "$0")/.."
python profiles/sweep.py "$@"
