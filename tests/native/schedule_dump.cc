// Test helper (CPU only): builds a net exactly as the engine does, runs plan_streams and prints one line per op with its
// stream, the ops it waits for and the memory it touches, for tests/test_schedule_cpu.py to check independently.
//   schedule_dump <model> <max_streams>
#include <cstdio>
#include <cstdlib>
#include <string>

#include "graph.h"

static void acc(const dml::Net& net, int t, bool blocks) {
  const dml::Tensor& tt = net.tensors[t];
  int c1 = tt.c_off + tt.c;
  if (blocks) {
    c1 = tt.c_off + (tt.c + 63) / 64 * 64;
    if (c1 > tt.c_pitch) c1 = tt.c_pitch;
  }
  long long a0 = 0, a1 = 0;
  if (tt.buf >= 0) {
    a0 = net.bufs[tt.buf].offset;
    a1 = a0 + (net.bufs[tt.buf].elems + 63) / 64 * 64;
  }
  printf(" %d:%d:%d:%lld:%lld", tt.buf, tt.c_off, c1, a0, a1);
}

int main(int argc, char** argv) {
  const int model = argc > 1 ? atoi(argv[1]) : 1;
  const int ns = argc > 2 ? atoi(argv[2]) : 4;
  dml::Net net;
  std::string err;
  if (!dml::build_net(model, false, 2, true, true, &net, &err, true)) {
    fprintf(stderr, "%s\n", err.c_str());
    return 1;
  }
  dml::plan_streams(&net, ns);
  printf("streams %d regions %zu arena %lld\n", net.n_streams, net.regions.size(), (long long)net.arena_elems);
  for (size_t i = 0; i < net.ops.size(); ++i) {
    const dml::Op& op = net.ops[i];
    printf("op %zu %s skip=%d stream=%d region=%d solo=%d signal=%d waits=", i, op.name.c_str(), op.skip ? 1 : 0, op.stream, op.region,
           op.solo ? 1 : 0, op.signal ? 1 : 0);
    for (size_t w = 0; w < op.waits.size(); ++w) printf("%s%d", w ? "," : "", op.waits[w]);
    printf(" R");
    acc(net, op.in, op.kind == dml::OP_CONV);
    if (op.res >= 0) acc(net, op.res, false);
    if (op.in2 >= 0) acc(net, op.in2, true);
    printf(" W");
    acc(net, op.merged_out >= 0 ? op.merged_out : op.out, false);
    printf("\n");
  }
  return 0;
}
