#!/bin/bash
# integer-pipe issue rates (tools/microbench.cu) + the mnemonics really present in the binary
mkdir -p gpurun_out
tools/microbench > gpurun_out/r02_microbench_int_pipes.jsonl
cat gpurun_out/r02_microbench_int_pipes.jsonl
cuobjdump -sass tools/microbench | grep -oE "^\s+/\*[0-9a-f]+\*/\s+[A-Za-z0-9_.]+" | awk '{print $2}' | sort | uniq -c | sort -rn | head -16 > gpurun_out/r02_microbench_sass_counts.txt
