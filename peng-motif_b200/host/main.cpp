// main.cpp — `peng_motif SEQFILE [options]`: the reference's command line (src/main.cpp) driving the B200 hot path.
#include "peng_host.h"

int main(int argc, char** argv) {
  const peng_host::Options options = peng_host::parse_command_line(argc, argv);
  return peng_host::run(options);
}
