#!/bin/bash
# host topology of the GPU box (for the end-to-end scaling numbers): CPUs, NUMA nodes, GPU <-> CPU affinity
nproc; cat /sys/fs/cgroup/cpu.max 2>/dev/null; lscpu | grep -i "socket\|numa\|model name\|^cpu(s)"; 
nvidia-smi topo -m 2>&1 | head -30
for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/vendor 2>/dev/null)" = "0x10de" ] && [ "$(cat $d/class 2>/dev/null | cut -c1-4)" = "0x03" ]; then echo "$d numa=$(cat $d/numa_node) link=$(cat $d/current_link_speed 2>/dev/null) x$(cat $d/current_link_width 2>/dev/null)"; fi; done
python - <<'PY'
import os; print("affinity", sorted(os.sched_getaffinity(0)))
PY
free -g | head -2
