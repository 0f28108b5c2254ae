// build/show_example: print the first -n examples of a RecordIO file of Example protos
// (ref src/data/show_example.h):  build/show_example -input train.recordio -n 3
#include "data/example.h"
#include "util/file.h"
#include "util/flags.h"

namespace PS {
DECLARE_string(input);  // data/data_flags.cc
DEFINE_int32(n, 3, "number of examples to print");
}  // namespace PS

int main(int argc, char* argv[]) {
  using namespace PS;
  string err;
  if (!FlagRegistry::get().parse(&argc, &argv, true, &err) || FLAGS_input == "stdin") {
    fprintf(stderr, "%s\nusage: show_example -input FILE.recordio [-n 3]\n", err.c_str());
    return 1;
  }
  File* in = File::openOrDie(FLAGS_input, "r");
  RecordReader reader(in);
  Example ex;
  for (int i = 0; i < FLAGS_n && reader.ReadProtocolMessage(&ex); ++i) {
    printf("example %d\n", i);
    for (const Slot& s : ex.slot) {
      printf("  slot { id: %d", s.id);
      for (size_t k = 0; k < s.key.size(); ++k) printf(" key: %llu", (unsigned long long)s.key[k]);
      for (size_t k = 0; k < s.val.size(); ++k) printf(" val: %g", s.val[k]);
      printf(" }\n");
    }
  }
  in->close();
  return 0;
}
