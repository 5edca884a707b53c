#!/bin/bash
# host_topology.sh -- what the box says about host locality: NUMA nodes, the CPUs / memory nodes this container may
# use, and the node + local CPUs of every GPU (input to mvs_bind_host_to_device).  Prints to stdout.
echo "nproc: $(nproc)   online cpus: $(cat /sys/devices/system/cpu/online)   numa nodes: $(cat /sys/devices/system/node/online 2>/dev/null)"
grep -E "Cpus_allowed_list|Mems_allowed_list" /proc/self/status
for n in /sys/devices/system/node/node*; do
    [ -d "$n" ] && echo "$(basename $n): cpus $(cat $n/cpulist)  $(grep MemTotal $n/meminfo | awk '{print $4, $5}')"
done
nvidia-smi --query-gpu=index,pci.bus_id --format=csv,noheader | while IFS=', ' read idx bus; do
    b=$(echo "${bus#0000}" | tr 'A-F' 'a-f')
    for cand in /sys/bus/pci/devices/*"$b"; do
        [ -e "$cand/numa_node" ] && echo "gpu $idx $bus numa_node=$(cat $cand/numa_node) local_cpulist=$(cat $cand/local_cpulist)"
    done
done
nvidia-smi topo -m 2>&1 | head -30
