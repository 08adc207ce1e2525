// hostnuma.cpp -- see hostnuma.hpp
#include "hostnuma.hpp"
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

namespace ccbh {

static bool read_line(const char* path, char* buf, size_t cap)
{
    FILE* f = fopen(path, "r");
    if (!f) return false;
    const bool ok = fgets(buf, (int)cap, f) != nullptr;
    fclose(f);
    return ok;
}

int pci_numa_node(const char* pci_bus_id)
{
    if (!pci_bus_id || !*pci_bus_id) return -1;
    char id[64]; size_t n = 0;
    for (const char* p = pci_bus_id; *p && n + 1 < sizeof id; p++) id[n++] = (char)tolower((unsigned char)*p);
    id[n] = 0;
    char path[160], line[64];
    snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", id);
    if (!read_line(path, line, sizeof line)) return -1;
    return atoi(line);                                   // -1 on a single-node box
}

// "0-31,64-95\n"
bool node_cpus(int node, cpu_set_t* out)
{
    char path[96], line[4096];
    snprintf(path, sizeof path, "/sys/devices/system/node/node%d/cpulist", node);
    if (!read_line(path, line, sizeof line)) return false;
    CPU_ZERO(out);
    int any = 0;
    for (char* p = line; *p && *p != '\n';) {
        char* e;
        long a = strtol(p, &e, 10), b = a;
        if (e == p) return false;
        if (*e == '-') { p = e + 1; b = strtol(p, &e, 10); if (e == p) return false; }
        for (long c = a; c <= b && c < CPU_SETSIZE; c++) { CPU_SET((int)c, out); any = 1; }
        p = *e == ',' ? e + 1 : e;
        if (*e && *e != ',' && *e != '\n') return false;
    }
    return any != 0;
}

NodeBind::NodeBind(const char* pci_bus_id)
{
    const char* env = getenv("CCB_NUMA");
    if (env && atoi(env) == 0) return;
    const int node = pci_numa_node(pci_bus_id);
    if (node < 0) return;
    cpu_set_t want, now;
    if (!node_cpus(node, &want)) return;
    if (sched_getaffinity(0, sizeof old_, &old_) != 0) return;
    CPU_AND(&now, &want, &old_);                         // stay inside what the thread was allowed before
    if (CPU_COUNT(&now) == 0) return;
    if (sched_setaffinity(0, sizeof now, &now) != 0) return;
    node_ = node;
}

NodeBind::~NodeBind()
{
    if (node_ >= 0) sched_setaffinity(0, sizeof old_, &old_);
}

}  // namespace ccbh
