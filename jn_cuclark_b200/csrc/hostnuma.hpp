// hostnuma.hpp -- pinned batch buffers on the memory node next to their GPU.
//
// A two-socket box hangs half of its GPUs off each socket.  Pinned memory is placed on the node of the CPU the
// allocating thread runs on, so a process (or an OpenMP worker) that happens to run on the far socket gets buffers
// whose every DMA crosses the socket interconnect -- with eight GPUs fed at once that link saturates first.
// NodeBind moves the CALLING THREAD onto the CPUs of the device's node for its lifetime (sysfs: PCI device ->
// numa_node -> cpulist) and puts the old affinity back afterwards; the pages a cudaMallocHost inside that window pins
// are local to the GPU that will read them.  Does nothing when the box has one node, when sysfs does not say, when the
// thread's affinity mask excludes the whole node (a container's cpuset), or under CCB_NUMA=0.
#pragma once
#include <sched.h>

namespace ccbh {

// node of a PCI device ("0000:1b:00.0", any case), -1 = unknown / single node
int pci_numa_node(const char* pci_bus_id);
// CPUs of a node out of /sys/devices/system/node/node<N>/cpulist; false = unknown
bool node_cpus(int node, cpu_set_t* out);

class NodeBind {
public:
    explicit NodeBind(const char* pci_bus_id);
    ~NodeBind();
    NodeBind(const NodeBind&) = delete;
    NodeBind& operator=(const NodeBind&) = delete;
    int node() const { return node_; }          // node bound to, -1 = nothing changed
private:
    cpu_set_t old_;
    int node_ = -1;
};

}  // namespace ccbh
