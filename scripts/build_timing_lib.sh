# Development build of the library with per-phase clocks in peel_flow_kernel -> build/libevoxb_timing.so
set -e
cd "$(dirname "$0")/.."
mkdir -p build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -shared -cudart static -ldl \
     -DEVOXB_FLOW_TIMING=1 -o build/libevoxb_timing.so evox-myc_b200/csrc/evoxb.cu
