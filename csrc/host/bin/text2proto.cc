// build/text2proto: text examples -> RecordIO of binary Example protos + an ExampleInfo text file.
//   build/text2proto -input train.txt -format libsvm -output train      (writes train.recordio, train.info)
// The reference sketches this tool in src/data/text2proto.h (its body is commented out); the
// output is what a DataConfig { format: PROTO } reads.
#include <algorithm>

#include "data/text_parser.h"
#include "util/file.h"

namespace PS {
DECLARE_string(input);   // data/data_flags.cc
DECLARE_string(output);
DECLARE_string(format);
}  // namespace PS

int main(int argc, char* argv[]) {
  using namespace PS;
  string err;
  if (!FlagRegistry::get().parse(&argc, &argv, true, &err) || FLAGS_input == "stdin" || FLAGS_output.empty()) {
    fprintf(stderr, "%s\nusage: text2proto -input FILE -format libsvm -output PREFIX\n", err.c_str());
    return 1;
  }
  string name = FLAGS_format == "none" ? "libsvm" : FLAGS_format;
  if (name.compare(0, 3, "ps_") == 0) name = name.substr(3);
  std::transform(name.begin(), name.end(), name.begin(), ::toupper);
  DataConfigView::TextFormat format;
  if (!DataConfigView::TextFormat_Parse(name, &format)) {
    fprintf(stderr, "unknown text format '%s'\n", FLAGS_format.c_str());
    return 1;
  }
  ExampleParser parser;
  parser.init(format);
  InfoParser info;
  File* out = File::openOrDie(FLAGS_output + ".recordio", "w");
  RecordWriter writer(out);
  Example ex;
  uint64 bad = 0, written = 0;
  Timer timer;
  timer.start();
  FileLineReader reader(FLAGS_input);
  reader.set_line_callback([&](char* line) {
    if (!*line) return;
    if (!parser.toProto(line, &ex) || !info.add(ex)) { ++bad; return; }
    CHECK(writer.WriteProtocolMessage(ex));
    ++written;
  });
  reader.Reload();
  out->close();
  if (!reader.loaded_successfully()) {
    fprintf(stderr, "cannot read %s\n", FLAGS_input.c_str());
    return 1;
  }
  CHECK(writeProtoToASCIIFile(info.info(), FLAGS_output + ".info"));
  timer.stop();
  fprintf(stderr, "written %llu examples in %.3f sec", (unsigned long long)written, timer.get());
  if (bad) fprintf(stderr, ", skipped %llu malformed lines", (unsigned long long)bad);
  fprintf(stderr, "\n");
  return 0;
}
