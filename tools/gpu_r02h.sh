#!/bin/bash
# tests + bench + tile phase profile, then the round's profile set (tag r02e)
bash tools/gpu_iter.sh
bash tools/gpu_profile.sh r02e
