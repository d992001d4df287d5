#!/bin/bash
# Host / PCIe topology of the GPU box (for the e2e feeder: which NUMA node each GPU hangs off).
mkdir -p gpurun_out
{
  nvidia-smi topo -m
  lscpu | head -40
  echo "affinity: $(python -c 'import os;print(len(os.sched_getaffinity(0)), sorted(os.sched_getaffinity(0))[:4])')"
  for d in /sys/devices/system/node/node*; do echo "$d $(cat $d/cpulist) $(grep MemTotal $d/meminfo)"; done
  nvidia-smi --query-gpu=index,pci.bus_id,pcie.link.gen.current,pcie.link.width.current --format=csv
  for p in /sys/bus/pci/devices/*; do c=$(cat $p/class 2>/dev/null); if [ "$c" = "0x030200" ]; then echo "$p numa=$(cat $p/numa_node)"; fi; done
} > gpurun_out/topo.txt 2>&1
