#!/bin/bash
# What the GPU box has (kept under profiles/): the toolchains that decide the CPU baseline, host
# cores and memory, the GPU topology. Read-only probes.
echo "== date";            date -u
echo "== go version";      (go version 2>&1 || echo "go: not found")
echo "== other toolchains"; (javac -version 2>&1 | head -1 || true); (node --version 2>&1 | head -1 || true); gcc --version | head -1; nvcc --version | tail -2
echo "== nproc";           nproc
echo "== lscpu";           lscpu | grep -E "Model name|Socket|Core|Thread|NUMA|MHz" | head -12
echo "== memory";          grep -E "MemTotal|MemAvailable" /proc/meminfo
echo "== nvidia-smi";      nvidia-smi --query-gpu=index,name,memory.total,clocks.max.sm,clocks.max.mem,pcie.link.gen.max,pcie.link.width.max --format=csv
echo "== topology";        nvidia-smi topo -m 2>&1 | head -30
echo "== zlib / tar / gzip"; python -c "import zlib; print('zlib', zlib.ZLIB_VERSION, zlib.ZLIB_RUNTIME_VERSION)"; tar --version | head -1; gzip --version | head -1
