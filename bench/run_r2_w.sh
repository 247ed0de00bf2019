#!/bin/bash
cd "$(dirname "$0")/.."
python bench/probe_tma.py 2>&1 | tail -8
