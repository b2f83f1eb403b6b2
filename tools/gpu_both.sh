#!/bin/bash
bash tools/gpu_quick.sh
bash tools/gpu_variants.sh
