#!/bin/bash
python tools/dials_timeline.py 2>&1 | tail -14
