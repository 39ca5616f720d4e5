#!/bin/bash
bash tools/gpu_l.sh
bash tools/gpu_m.sh 2>&1 | grep "gpu__time_duration"
