// flags of the data tools (ref src/data/common.cc:7-9, src/data/text_parser.cc:9,
// src/util/filelinereader.cc:13)
#include "util/flags.h"

namespace PS {
DEFINE_string(input, "stdin", "input file name (text2proto / show_example tools)");
DEFINE_string(output, "", "output file name");
DEFINE_string(format, "none", "text format: libsvm, adfea, terafea, ps_sparse, ps_sparse_binary, ps_dense");
DEFINE_bool(shuffle_fea_id, false, "hash the feature ids of TERAFEA data with MurmurHash3");
DEFINE_int32(line_limit, 0, "maximal number of lines to read, 0 = all");
}  // namespace PS
