// Native core of the schedule compiler (compile/plan.py, compile/list_schedule.py): plan simulation and the
// memory-bounded gather hoisting loop.
//
// Reference counterpart: the C++ runtime behind `DeepCompileBuilder` (`compile/util.py`, `compile/init_z3.py:24-49`,
// `compile/list_schedule.py:171-209`) — absent from the reference tree.  The Python implementation stays as the
// oracle; this one exists because hoisting re-simulates the plan after every candidate swap (O(N^2) simulations of
// O(N) nodes), which is seconds in Python for models with hundreds of units and microseconds here.
//
// A plan is passed as parallel arrays, one entry per node:
//   kind  : 0 gather, 1 compute, 2 release, 3 reduce, 4 offload, 5 reload
//   unit  : unit index (-1 = not tied to a unit)
//   phase : 0 forward, 1 backward
//   bytes : memory acquired / freed
//   time  : microseconds
#include <algorithm>
#include <cstdint>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#define SXBH_API extern "C" __attribute__((visibility("default")))

namespace {

enum Kind { kGather = 0, kCompute = 1, kRelease = 2, kReduce = 3, kOffload = 4, kReload = 5 };

struct Node {
  int kind, unit, phase;
  long long bytes;
  double time_us;
};

struct Sim {
  double peak_bytes, time_us, exposed_us;
};

// Mirrors ExecutionPlan.simulate: gathered-bytes profile + two-stream time estimate.
Sim simulate(const std::vector<Node>& nodes, const std::unordered_set<int>& persistent, long long base_bytes) {
  long long live = base_bytes, peak = base_bytes;
  std::unordered_map<int, bool> gathered;
  for (int u : persistent) gathered[u] = true;
  std::unordered_map<int, double> ready;
  double t_comm = 0.0, t_comp = 0.0, exposed = 0.0;
  for (const Node& n : nodes) {
    switch (n.kind) {
      case kGather:
      case kReload: {
        if (n.kind == kGather) {
          auto it = gathered.find(n.unit);
          if (it != gathered.end() && it->second) break;
        }
        live += n.bytes;
        peak = std::max(peak, live);
        gathered[n.unit] = true;
        t_comm = std::max(t_comm, t_comp) + n.time_us;
        ready[n.unit] = t_comm;
        break;
      }
      case kRelease:
      case kOffload: {
        if (n.kind == kRelease) {
          auto it = gathered.find(n.unit);
          if (persistent.count(n.unit) || it == gathered.end() || !it->second) break;
        }
        live -= n.bytes;
        gathered[n.unit] = false;
        break;
      }
      case kCompute: {
        auto it = ready.find(n.unit);
        const double need = it == ready.end() ? 0.0 : it->second;
        if (need > t_comp) {
          exposed += need - t_comp;
          t_comp = need;
        }
        t_comp += n.time_us;
        break;
      }
      case kReduce:
        t_comm = std::max(t_comm, t_comp) + n.time_us;
        break;
      default:
        break;
    }
  }
  return {static_cast<double>(peak), std::max(t_comp, t_comm), exposed};
}

std::vector<Node> pack(int n, const int* kind, const int* unit, const int* phase, const long long* bytes,
                       const double* time_us) {
  std::vector<Node> v(n);
  for (int i = 0; i < n; ++i) v[i] = {kind[i], unit[i], phase[i], bytes[i], time_us[i]};
  return v;
}

}  // namespace

// out[0] = peak bytes, out[1] = estimated time (us), out[2] = exposed communication (us).
SXBH_API int sxbh_plan_simulate(int n, const int* kind, const int* unit, const int* phase, const long long* bytes,
                                const double* time_us, const int* persistent, int n_persistent,
                                long long base_bytes, double* out) {
  if (n < 0 || !out) return -1;
  std::unordered_set<int> pers(persistent, persistent + n_persistent);
  const Sim s = simulate(pack(n, kind, unit, phase, bytes, time_us), pers, base_bytes);
  out[0] = s.peak_bytes; out[1] = s.time_us; out[2] = s.exposed_us;
  return 0;
}

// Hoist every non-persistent GATHER towards the front, one position at a time, while the simulated peak stays within
// `budget`; never across another gather, across the RELEASE of the same unit, or across a non-compute node of the other
// phase.  `perm[i]` = index (in the input arrays) of the node that ends up at position i.
SXBH_API int sxbh_plan_hoist_gathers(int n, const int* kind, const int* unit, const int* phase,
                                     const long long* bytes, const double* time_us, const int* persistent,
                                     int n_persistent, long long budget, int* perm) {
  if (n < 0 || !perm) return -1;
  std::unordered_set<int> pers(persistent, persistent + n_persistent);
  std::vector<Node> nodes = pack(n, kind, unit, phase, bytes, time_us);
  std::vector<int> idx(n);
  for (int i = 0; i < n; ++i) idx[i] = i;
  for (int i = 0; i < n; ++i) {
    const Node cur = nodes[i];
    if (cur.kind != kGather || pers.count(cur.unit)) continue;
    int pos = i;
    while (pos > 0) {
      const Node& prev = nodes[pos - 1];
      if (prev.phase != cur.phase && prev.kind != kCompute) break;
      if (prev.kind == kGather) break;
      if (prev.kind == kRelease && prev.unit == cur.unit) break;
      std::swap(nodes[pos - 1], nodes[pos]);
      std::swap(idx[pos - 1], idx[pos]);
      if (simulate(nodes, pers, 0).peak_bytes > static_cast<double>(budget)) {
        std::swap(nodes[pos - 1], nodes[pos]);
        std::swap(idx[pos - 1], idx[pos]);
        break;
      }
      --pos;
    }
  }
  for (int i = 0; i < n; ++i) perm[i] = idx[i];
  return 0;
}
