#!/bin/bash
# prints the host topology facts that bound multi-GPU end-to-end throughput
echo "cpus allowed: $(grep Cpus_allowed_list /proc/self/status)"; echo "mems allowed: $(grep Mems_allowed_list /proc/self/status)"
ls -d /sys/devices/system/node/node* 2>/dev/null | tr '\n' ' '; echo
for n in /sys/devices/system/node/node*; do echo "$n cpus $(cat $n/cpulist) mem $(grep MemTotal $n/meminfo | awk '{print $4,$5}')"; done
nvidia-smi --query-gpu=index,pci.bus_id --format=csv,noheader | while IFS=, read idx bus; do b=$(echo $bus | tr 'A-Z' 'a-z' | sed 's/^ *0000//'); echo "gpu $idx $bus numa $(cat /sys/bus/pci/devices/$b/numa_node 2>/dev/null) local_cpus $(cat /sys/bus/pci/devices/$b/local_cpulist 2>/dev/null)"; done
nvidia-smi topo -m 2>/dev/null | head -14
lscpu | grep -i "model name\|socket\|numa\|^cpu(s)"
