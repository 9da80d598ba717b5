python -m pytest tests -m gpu -x -q > gpurun_out/s32_pytest.log 2>&1; tail -3 gpurun_out/s32_pytest.log
for g in 50 159 500; do for ml in 1 2 4 8; do echo "g=$g max_leaf=$ml"; python tools/prof_render.py 8 f32 2 1 0 $g 3840 $ml | tail -1; done; done > gpurun_out/s32_leaf.log 2>&1
cat gpurun_out/s32_leaf.log
