// Typed views of every message the reference defines in its .proto files, over the dynamic
// text-format tree of util/textproto.h.  Field names, enum names, defaults and extension
// names follow the schemas so that the shipped *.conf files parse unchanged:
//   ref src/system/proto/{task,node,heartbeat}.proto, src/parameter/proto/param.proto,
//       src/filter/proto/filter.proto, src/data/proto/{data,example}.proto,
//       src/util/proto/{range,matrix,auc}.proto, src/learner/proto/{bcd,sgd}.proto,
//       src/app/linear_method/proto/linear.proto, src/app/main/proto/app.proto
// `Example` / `Slot` are the exception: they are bulk data, kept as plain structs with a
// protobuf *binary* codec (data/example.h) so RecordIO files stay compatible.
#pragma once
#include "util/range.h"
#include "util/textproto.h"

namespace PS {

#define PB_DEFINE_ENUM_IN_CLASS(Enum, LIST)                                      \
  enum Enum { LIST(PB_ENUM_ITEM) };                                              \
  static const char* Enum##_Name(Enum v) {                                       \
    switch (v) { LIST(PB_ENUM_CASE) }                                            \
    return "?";                                                                  \
  }                                                                              \
  static bool Enum##_Parse(const std::string& s, Enum* out) {                    \
    LIST(PB_ENUM_PARSE)                                                          \
    if (!s.empty() && (isdigit((unsigned char)s[0]) || s[0] == '-')) {           \
      *out = (Enum)atoi(s.c_str());                                              \
      return true;                                                               \
    }                                                                            \
    return false;                                                                \
  }

// ---- util/proto/range.proto ------------------------------------------------------------------
class PbRangeView : public PbView {
 public:
  PB_VIEW(PbRangeView)
  PB_UINT64(begin, 0)
  PB_UINT64(end, 0)
  template <typename T> void from(const Range<T>& r) { set_begin((uint64)r.begin()); set_end((uint64)r.end()); }
  template <typename T> Range<T> to() const { return Range<T>((T)begin(), (T)end()); }
};
typedef Pb<PbRangeView> PbRange;
template <typename T> inline Range<T> toRange(const PbRangeView& v) { return v.to<T>(); }

// ---- data type tags (task.proto DataType) ----------------------------------------------------
#define PS_DATATYPE_LIST(X) X(OTHER, 0) X(INT8, 1) X(INT16, 2) X(INT32, 3) X(INT64, 4) \
  X(UINT8, 5) X(UINT16, 6) X(UINT32, 7) X(UINT64, 8) X(FLOAT, 9) X(DOUBLE, 10) X(CHAR, 11)
PB_DEFINE_ENUM(DataType, PS_DATATYPE_LIST)

template <typename V> inline DataType encodeType() { return OTHER; }
template <> inline DataType encodeType<int8>() { return INT8; }
template <> inline DataType encodeType<int16>() { return INT16; }
template <> inline DataType encodeType<int32>() { return INT32; }
template <> inline DataType encodeType<int64>() { return INT64; }
template <> inline DataType encodeType<uint8>() { return UINT8; }
template <> inline DataType encodeType<uint16>() { return UINT16; }
template <> inline DataType encodeType<uint32>() { return UINT32; }
template <> inline DataType encodeType<uint64>() { return UINT64; }
template <> inline DataType encodeType<float>() { return FLOAT; }
template <> inline DataType encodeType<double>() { return DOUBLE; }
template <> inline DataType encodeType<char>() { return CHAR; }
inline size_t sizeofType(DataType t) {
  switch (t) {
    case INT8: case UINT8: case CHAR: return 1;
    case INT16: case UINT16: return 2;
    case INT32: case UINT32: case FLOAT: return 4;
    case INT64: case UINT64: case DOUBLE: return 8;
    default: return 1;
  }
}

// ---- param.proto ------------------------------------------------------------------------------
#define PS_OPERATOR_LIST(X) X(PLUS, 0) X(MINUS, 1) X(TIMES, 2) X(DIVIDE, 3) X(AND, 4) X(OR, 5) X(XOR, 6)
PB_DEFINE_ENUM(Operator, PS_OPERATOR_LIST)

class ParameterInitConfigView : public PbView {
 public:
  PB_VIEW(ParameterInitConfigView)
#define L(X) X(ZERO, 1) X(CONSTANT, 2) X(GAUSSIAN, 3) X(FILE, 4) X(CLONE, 5)
  PB_DEFINE_ENUM_IN_CLASS(Type, L)
#undef L
  PB_ENUM(Type, type, ZERO)
  PB_DOUBLE(constant, 1)
  PB_DOUBLE(mean, 0)
  PB_DOUBLE(std, 1)
  PB_STRING(file_name, "")
};
typedef Pb<ParameterInitConfigView> ParameterInitConfig;

class TimestampView : public PbView {
 public:
  PB_VIEW(TimestampView)
  PB_STRING(sender, "")
  PB_INT32(time, 0)
};

class TailKeyFilterView : public PbView {
 public:
  PB_VIEW(TailKeyFilterView)
  PB_BOOL(insert_count, false)
  PB_INT32(query_key, 0)
  PB_BOOL(query_value, false)
  PB_INT32(countmin_n, 0)
  PB_INT32(countmin_k, 0)
};

class CallSharedParaView : public PbView {
 public:
  PB_VIEW(CallSharedParaView)
#define L(X) X(PUSH, 1) X(PULL, 2)
  PB_DEFINE_ENUM_IN_CLASS(Command, L)
#undef L
  PB_ENUM(Command, cmd, PUSH)
  PB_ENUM(Operator, op, PLUS)
  PB_MSG(TailKeyFilterView, tail_filter)
  PB_BOOL(insert_key, false)
  PB_BOOL(gather, false)
  PB_BOOL(replica, false)
  PB_REPEATED_MSG(TimestampView, backup)
};
typedef Pb<CallSharedParaView> CallSharedPara;

// ---- filter.proto -----------------------------------------------------------------------------
class FixedFloatConfigView : public PbView {
 public:
  PB_VIEW(FixedFloatConfigView)
  PB_DOUBLE(min_value, -1)
  PB_DOUBLE(max_value, 1)
  PB_INT32(num_bytes, 3)
};

class FilterConfigView : public PbView {
 public:
  PB_VIEW(FilterConfigView)
#define L(X) X(KEY_CACHING, 1) X(COMPRESSING, 2) X(FIXING_FLOAT, 3)
  PB_DEFINE_ENUM_IN_CLASS(Type, L)
#undef L
  typedef FixedFloatConfigView FixedFloatConfig;
  PB_ENUM(Type, type, KEY_CACHING)
  PB_BOOL(clear_cache_if_done, false)
  PB_REPEATED_MSG(FixedFloatConfigView, fixed_point)
  PB_UINT64(signature, 0)
  PB_REPEATED_UINT64(uncompressed_size)
};
typedef Pb<FilterConfigView> FilterConfig;

// ---- data.proto -------------------------------------------------------------------------------
class HDFSConfigView : public PbView {
 public:
  PB_VIEW(HDFSConfigView)
  PB_STRING(home, "")
  PB_STRING(ugi, "")
  PB_STRING(namenode, "")
};

class DataConfigView : public PbView {
 public:
  PB_VIEW(DataConfigView)
#define L(X) X(BIN, 1) X(PROTO, 2) X(TEXT, 3)
  PB_DEFINE_ENUM_IN_CLASS(DataFormat, L)
#undef L
#define L(X) X(DENSE, 1) X(SPARSE, 2) X(SPARSE_BINARY, 3) X(ADFEA, 4) X(LIBSVM, 5) X(TERAFEA, 6) X(VW, 7)
  PB_DEFINE_ENUM_IN_CLASS(TextFormat, L)
#undef L
  PB_ENUM(DataFormat, format, TEXT)
  PB_ENUM(TextFormat, text, LIBSVM)
  PB_REPEATED_STRING(file)
  PB_MSG(HDFSConfigView, hdfs)
  PB_MSG(PbRangeView, range)
  PB_BOOL(ignore_feature_group, false)
  PB_INT32(max_num_files_per_worker, -1)
  PB_INT32(max_num_lines_per_file, -1)
};
typedef Pb<DataConfigView> DataConfig;

// ---- example.proto (metadata part) ------------------------------------------------------------
class SlotInfoView : public PbView {
 public:
  PB_VIEW(SlotInfoView)
#define L(X) X(DENSE, 1) X(SPARSE, 2) X(SPARSE_BINARY, 3)
  PB_DEFINE_ENUM_IN_CLASS(Format, L)
#undef L
  PB_ENUM(Format, format, SPARSE)
  PB_INT32(id, 0)
  PB_UINT64(min_key, 0xFFFFFFFFFFFFFFFFULL)
  PB_UINT64(max_key, 0)
  PB_UINT64(nnz_ele, 0)
  PB_UINT64(nnz_ex, 0)
};
typedef Pb<SlotInfoView> SlotInfo;

class ExampleInfoView : public PbView {
 public:
  PB_VIEW(ExampleInfoView)
  PB_REPEATED_MSG(SlotInfoView, slot)
  PB_UINT64(num_ex, 0)
};
typedef Pb<ExampleInfoView> ExampleInfo;

// ---- matrix.proto / auc.proto -----------------------------------------------------------------
class MatrixInfoView : public PbView {
 public:
  PB_VIEW(MatrixInfoView)
#define L(X) X(DENSE, 1) X(SPARSE, 2) X(SPARSE_BINARY, 3)
  PB_DEFINE_ENUM_IN_CLASS(Type, L)
#undef L
  PB_ENUM(Type, type, SPARSE)
  PB_BOOL(row_major, true)
  PB_INT32(id, 0)
  PB_MSG(PbRangeView, row)
  PB_MSG(PbRangeView, col)
  PB_UINT64(nnz, 0)
  PB_UINT64(sizeof_index, 0)
  PB_UINT64(sizeof_value, 0)
};
typedef Pb<MatrixInfoView> MatrixInfo;

class AUCDataView : public PbView {
 public:
  PB_VIEW(AUCDataView)
  PB_REPEATED_NUM(int64, getInt, addInt, tp_key)
  PB_REPEATED_UINT64(tp_count)
  PB_REPEATED_NUM(int64, getInt, addInt, fp_key)
  PB_REPEATED_UINT64(fp_count)
};
typedef Pb<AUCDataView> AUCData;

// ---- node.proto / heartbeat.proto -------------------------------------------------------------
class NodeView : public PbView {
 public:
  PB_VIEW(NodeView)
#define L(X) X(SERVER, 0) X(WORKER, 1) X(MANAGER, 2) X(SCHEDULER, 3) X(GROUP, 4) X(UNUSED, 5)
  PB_DEFINE_ENUM_IN_CLASS(Role, L)
#undef L
  PB_ENUM(Role, role, SERVER)
  PB_STRING(id, "")
  PB_INT32(rank, 0)
  PB_STRING(hostname, "")
  PB_INT32(port, 0)
  PB_MSG(PbRangeView, key)
};
typedef Pb<NodeView> Node;
typedef string NodeID;

class HeartbeatReportView : public PbView {
 public:
  PB_VIEW(HeartbeatReportView)
  PB_INT32(task_id, 0)
  PB_STRING(hostname, "")
  PB_UINT64(seconds_since_epoch, 0)
  PB_UINT64(total_time_milli, 0)
  PB_UINT64(busy_time_milli, 0)
  PB_UINT64(net_in_mb, 0)
  PB_UINT64(net_out_mb, 0)
  PB_UINT64(process_cpu_usage, 0)
  PB_UINT64(host_cpu_usage, 0)
  PB_UINT64(process_rss_mb, 0)
  PB_UINT64(process_virt_mb, 0)
  PB_UINT64(host_in_use_gb, 0)
  PB_UINT64(host_in_use_percentage, 0)
  PB_UINT64(host_net_in_bw, 0)
  PB_UINT64(host_net_out_bw, 0)
};
typedef Pb<HeartbeatReportView> HeartbeatReport;

// ---- bcd.proto --------------------------------------------------------------------------------
class BCDConfigView : public PbView {
 public:
  PB_VIEW(BCDConfigView)
  PB_DOUBLE(feature_block_ratio, 4)
  PB_BOOL(random_feature_block_order, true)
  PB_REPEATED_INT32(prior_fea_group)
  PB_INT32(num_iter_for_prior_fea_group, 5)
  PB_INT32(max_block_delay, 0)
  PB_INT32(max_pass_of_data, 10)
  PB_DOUBLE(epsilon, 1e-4)
  PB_INT32(tail_feature_freq, 0)
  PB_INT32(countmin_k, 2)
  PB_DOUBLE(countmin_n_ratio, 2.0)
  PB_INT32(max_num_parallel_groups_in_preprocessing, 1000)
  PB_INT32(max_data_buf_size_in_mb, 1000)
  PB_MSG(DataConfigView, local_cache)
  PB_MSG(ParameterInitConfigView, init_w)
  // extensions declared in linear.proto:22-32
  PB_EXT_DOUBLE(delta_init_value, "PS.LM.delta_init_value", 1)
  PB_EXT_DOUBLE(delta_max_value, "PS.LM.delta_max_value", 5)
  PB_EXT_DOUBLE(kkt_filter_threshold_ratio, "PS.LM.kkt_filter_threshold_ratio", 10)
};
typedef Pb<BCDConfigView> BCDConfig;

class BCDProgressView : public PbView {
 public:
  PB_VIEW(BCDProgressView)
  PB_DOUBLE(objective, 0)
  PB_DOUBLE(relative_obj, 0)
  PB_UINT64(nnz_w, 0)
  PB_DOUBLE(violation, 0)
  PB_UINT64(nnz_active_set, 0)
  PB_DOUBLE(total_time, 0)
  PB_REPEATED_DOUBLE(busy_time)
};
typedef Pb<BCDProgressView> BCDProgress;

class BCDCallView : public PbView {
 public:
  PB_VIEW(BCDCallView)
#define L(X) X(LOAD_DATA, 1) X(PREPROCESS_DATA, 2) X(UPDATE_MODEL, 3) X(EVALUATE_PROGRESS, 4) \
  X(SAVE_MODEL, 5) X(RECOVER, 6) X(COMPUTE_VALIDATION_AUC, 7)
  PB_DEFINE_ENUM_IN_CLASS(Command, L)
#undef L
  PB_ENUM(Command, cmd, LOAD_DATA)
  PB_MSG(PbRangeView, key)
  PB_DOUBLE(kkt_filter_threshold, 0)
  PB_BOOL(reset_kkt_filter, false)
  PB_REPEATED_INT32(fea_grp)
  PB_BOOL(hit_cache, false)
  PB_MSG(DataConfigView, data)
};
typedef Pb<BCDCallView> BCDCall;

class LoadDataReturnView : public PbView {
 public:
  PB_VIEW(LoadDataReturnView)
  PB_MSG(ExampleInfoView, example_info)
  PB_INT32(hit_cache, 0)
};
typedef Pb<LoadDataReturnView> LoadDataReturn;

// ---- sgd.proto --------------------------------------------------------------------------------
class SGDProgressView : public PbView {
 public:
  PB_VIEW(SGDProgressView)
  PB_REPEATED_DOUBLE(objective)
  PB_UINT64(num_examples_processed, 0)
  PB_REPEATED_DOUBLE(accuracy)
  PB_REPEATED_DOUBLE(auc)
  PB_UINT64(nnz, 0)
  PB_DOUBLE(weight_sum, 0)
  PB_DOUBLE(delta_sum, 0)
};
typedef Pb<SGDProgressView> SGDProgress;

class SGDCallView : public PbView {
 public:
  PB_VIEW(SGDCallView)
#define L(X) X(UPDATE_MODEL, 1) X(REPORT_PROGRESS, 2) X(SAVE_MODEL, 3) X(RECOVER, 4) X(COMPUTE_VALIDATION_AUC, 5)
  PB_DEFINE_ENUM_IN_CLASS(Command, L)
#undef L
  PB_ENUM(Command, cmd, UPDATE_MODEL)
  PB_MSG(DataConfigView, data)
  PB_INT32(report_interval, 1)
  PB_INT32(data_buf, 1000)
};
typedef Pb<SGDCallView> SGDCall;

// ---- task.proto -------------------------------------------------------------------------------
class ManageNodeView : public PbView {
 public:
  PB_VIEW(ManageNodeView)
#define L(X) X(CONNECT, 1) X(ADD, 2) X(UPDATE, 3) X(REPLACE, 4) X(REMOVE, 5)
  PB_DEFINE_ENUM_IN_CLASS(Command, L)
#undef L
  PB_ENUM(Command, cmd, CONNECT)
  PB_REPEATED_MSG(NodeView, node)
};
typedef Pb<ManageNodeView> ManageNode;

class ManageAppView : public PbView {
 public:
  PB_VIEW(ManageAppView)
#define L(X) X(ADD, 1) X(REMOVE, 2) X(RUN, 3) X(INIT, 4) X(DONE, 5)
  PB_DEFINE_ENUM_IN_CLASS(Command, L)
#undef L
  PB_ENUM(Command, cmd, ADD)
  PB_STRING(conf, "")
};
typedef Pb<ManageAppView> ManageApp;

class TaskView : public PbView {
 public:
  PB_VIEW(TaskView)
#define L(X) X(TERMINATE, 1) X(TERMINATE_CONFIRM, 2) X(REPLY, 3) X(MANAGE, 4) X(CALL_CUSTOMER, 5) X(HEARTBEATING, 6) X(PEER_CREDIT, 7)
  PB_DEFINE_ENUM_IN_CLASS(Type, L)
#undef L
  PB_ENUM(Type, type, CALL_CUSTOMER)
  PB_BOOL(request, false)
  PB_STRING(customer, "")
  PB_INT32(time, 0)
  PB_REPEATED_INT32(wait_time)
  PB_MSG(PbRangeView, key_range)
  PB_INT32(key_channel, 0)
  PB_BOOL(has_key, false)
  PB_STRING(key_channel_str, "")
  PB_ENUM(DataType, key_type, OTHER)
  PB_REPEATED_ENUM(DataType, value_type)
  PB_REPEATED_MSG(FilterConfigView, filter)
  PB_STRING(msg, "")
  PB_MSG(ManageNodeView, mng_node)
  PB_MSG(ManageAppView, mng_app)
  PB_MSG(CallSharedParaView, shared_para)
  PB_MSG(SGDCallView, sgd)
  PB_MSG(BCDCallView, bcd)
  // PeerVan (system/peer_van.h): value frames left in the sender's device arena
  // "index:alloc id:offset:length,...", and allocation ids the sender may reuse
  PB_STRING(peer_frames, "")
  PB_STRING(peer_credit, "")
};
typedef Pb<TaskView> Task;

// ---- linear.proto / app.proto -----------------------------------------------------------------
namespace LM {

class LossConfigView : public PbView {
 public:
  PB_VIEW(LossConfigView)
#define L(X) X(SQUARE, 1) X(LOGIT, 2) X(HINGE, 3) X(SQUARE_HINGE, 4)
  PB_DEFINE_ENUM_IN_CLASS(Type, L)
#undef L
  PB_ENUM(Type, type, LOGIT)
};
typedef Pb<LossConfigView> LossConfig;

class PenaltyConfigView : public PbView {
 public:
  PB_VIEW(PenaltyConfigView)
#define L(X) X(L1, 1) X(L2, 2)
  PB_DEFINE_ENUM_IN_CLASS(Type, L)
#undef L
  PB_ENUM(Type, type, L1)
  PB_REPEATED_DOUBLE(lambda)
};
typedef Pb<PenaltyConfigView> PenaltyConfig;

class LearningRateConfigView : public PbView {
 public:
  PB_VIEW(LearningRateConfigView)
#define L(X) X(CONSTANT, 1) X(DECAY, 2)
  PB_DEFINE_ENUM_IN_CLASS(Type, L)
#undef L
  PB_ENUM(Type, type, CONSTANT)
  PB_DOUBLE(alpha, 1)
  PB_DOUBLE(beta, 1)
};
typedef Pb<LearningRateConfigView> LearningRateConfig;

class SGDConfigView : public PbView {
 public:
  PB_VIEW(SGDConfigView)
#define L(X) X(STANDARD, 1) X(FTRL, 2)
  PB_DEFINE_ENUM_IN_CLASS(Algo, L)
#undef L
  PB_ENUM(Algo, algo, FTRL)
  PB_INT32(minibatch, 1000)
  PB_INT32(data_buf, 1000)
  PB_BOOL(ada_grad, true)
  PB_INT32(max_delay, 0)
  PB_INT32(num_data_pass, 1)
  PB_INT32(report_interval, 1)
  PB_INT32(tail_feature_freq, 0)
  PB_DOUBLE(countmin_n, 1e8)
  PB_INT32(countmin_k, 2)
  PB_INT32(fixing_float_by_nbytes, 0)
};
typedef Pb<SGDConfigView> SGDConfig;

class ConfigView : public PbView {
 public:
  PB_VIEW(ConfigView)
  PB_MSG(DataConfigView, training_data)
  PB_MSG(DataConfigView, validation_data)
  PB_MSG(DataConfigView, model_output)
  PB_MSG(DataConfigView, model_input)
  PB_MSG(LossConfigView, loss)
  PB_MSG(PenaltyConfigView, penalty)
  PB_MSG(LearningRateConfigView, learning_rate)
  PB_MSG(SGDConfigView, async_sgd)
  PB_MSG(BCDConfigView, darlin)
};
typedef Pb<ConfigView> Config;

}  // namespace LM

class AppConfigView : public PbView {
 public:
  PB_VIEW(AppConfigView)
  PB_MSG(LM::ConfigView, linear_method)
};
typedef Pb<AppConfigView> AppConfig;

}  // namespace PS
