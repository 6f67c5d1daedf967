# host topology of the GPU box as the ranks see it (NUMA nodes, cpuset, GPU affinity)
echo "== nodes online: $(cat /sys/devices/system/node/online 2>/dev/null)"
for n in /sys/devices/system/node/node*; do echo "$n cpus $(cat $n/cpulist) mem $(grep MemTotal $n/meminfo | awk '{print $4,$5}')"; done
echo "== allowed cpus: $(grep Cpus_allowed_list /proc/self/status)"; grep Mems_allowed_list /proc/self/status
nvidia-smi topo -m 2>&1 | head -30
for d in $(nvidia-smi --query-gpu=pci.bus_id --format=csv,noheader); do b=$(echo $d | tr 'A-Z' 'a-z' | sed 's/^0000//'); echo "$d numa_node $(cat /sys/bus/pci/devices/$b/numa_node 2>/dev/null)"; done
nproc
