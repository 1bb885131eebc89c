set -x
cd $GRAFT_REPO_ROOT
nproc
python tools/bench_pageable.py 1 2 3 4 6 8 > gpurun_out/r02h_pageable.jsonl 2> gpurun_out/r02h_pageable.err; tail -3 gpurun_out/r02h_pageable.err
cat gpurun_out/r02h_pageable.jsonl
python -m pytest tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -3
