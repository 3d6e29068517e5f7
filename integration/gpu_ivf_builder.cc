// GpuIVFBuilder — an IndexBuilder plugin class that trains, labels and packs an IVF index on the GPU and writes
// it in the reference's own on-disk format: the stock IVFSearcher and GpuIVFSearcher load the file.
//
// Mirrors IVFBuilder (src/core/algorithm/ivf/ivf_builder.{h,cc}) for its plain configuration — one level of
// centroids, fp32 rows, SquaredEuclidean or Cosine (CosineNormalizeConverter rows: d floats + the norm slot), no
// inverted-list quantizer, no optimizer quantizer, no original-feature segment:
//   train()  IVFBuilder::train(threads, holder) (ivf_builder.cc:209-270): a sample of the rows, Lloyd iterations
//            on the device (zvb200_kmeans = KmeansCluster::cluster), then the reference's IVFCentroidIndex is
//            built from the centroid list exactly as IVFBuilder::train(trainer) does (:272-330);
//   train(trainer)  the same entry the stock class has: centroids supplied by the caller;
//   build()  IVFBuilder::build -> build_label_index -> label (ivf_builder.cc:342-402, 602-649; ivf_builder.h:243-262):
//            one batched top-1 scan of all rows over the centroids (zvb200_ivf_assign), then the dump order
//            (zvb200_ivf_pack);
//   dump()   IVFBuilder::dump / dump_index (ivf_builder.cc:404-438, 651-721) through the reference's IVFDumper
//            (ivf_dumper.h:191-209), list by list, ids ascending — what the stock builder writes with one thread.
// Every other configuration is handed to an embedded stock IVFBuilder, whole (logged once).
#include <zvec/core/framework/index_builder.h>
#include <zvec/core/framework/index_cluster.h>
#include <zvec/core/framework/index_error.h>
#include <zvec/core/framework/index_factory.h>
#include <zvec/core/framework/index_helper.h>
#include <zvec/core/framework/index_logger.h>
#include <zvec/core/framework/index_meta.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "core/algorithm/ivf/ivf_centroid_index.h"
#include "core/algorithm/ivf/ivf_dumper.h"
#include "core/algorithm/ivf/ivf_params.h"

#include "../include/zvec_b200.h"

namespace zvec {
namespace core {

namespace {
const std::string kParamDevice("proxima.gpu.device");
const std::string kParamIterations("proxima.gpu.ivf.builder.max_iterations");  // Lloyd iterations (default 20, as KmeansCluster)
}  // namespace

class GpuIVFBuilder : public IndexBuilder {
 public:
  GpuIVFBuilder() = default;
  ~GpuIVFBuilder() override { this->cleanup(); }

  int init(const IndexMeta &meta, const ailego::Params &params) override {
    if (state_ != INIT) return IndexError_Logic;
    meta_ = meta;
    params_ = params;
    if (!IndexFactory::HasMetric(meta_.metric_name())) return IndexError_NoExist;
    if (!this->accelerated(meta, params)) {
      LOG_WARN("GpuIVFBuilder: this configuration is built by the stock IVFBuilder on the CPU");
      stock_ = IndexFactory::CreateBuilder("IVFBuilder");
      if (!stock_) return IndexError_NoExist;
      return stock_->init(meta, params);
    }
    // IVFBuilder::parse_centroids_num / parse_general_params (ivf_builder.cc:482-509, 535-580)
    const std::string cn = params.get_as_string(PARAM_IVF_BUILDER_CENTROID_COUNT);
    nlist_ = (uint32_t)std::strtoul(cn.c_str(), nullptr, 10);
    if (nlist_ == 0) return IndexError_InvalidArgument;
    sample_count_ = params.get_as_uint32(PARAM_IVF_BUILDER_TRAIN_SAMPLE_COUNT);
    sample_ratio_ = params.get_as_float(PARAM_IVF_BUILDER_TRAIN_SAMPLE_RATIO);
    params.get(PARAM_IVF_BUILDER_BLOCK_VECTOR_COUNT, &block_vector_count_);
    if (block_vector_count_ == 0) block_vector_count_ = 32;
    if (block_vector_count_ > 32 || (block_vector_count_ & (block_vector_count_ - 1))) return IndexError_InvalidArgument;
    if (block_vector_count_ * meta_.element_size() % 32 != 0) return IndexError_InvalidArgument;
    params.get(kParamDevice, &device_);
    params.get(kParamIterations, &max_iterations_);
    if (max_iterations_ == 0) max_iterations_ = 20;
    // the metas IVFBuilder::init derives (ivf_builder.cc:124-131)
    inner_meta_ = meta_;
    inner_meta_.set_reformer(std::string(), 0, ailego::Params());
    inner_meta_.set_converter(std::string(), 0, ailego::Params());
    cosine_ = meta_.metric_name() == "Cosine";
    state_ = INITED;
    return 0;
  }

  int cleanup(void) override {
    if (stock_) { stock_->cleanup(); stock_.reset(); }
    state_ = INIT;
    centroid_index_.reset();
    centroids_.clear(); rows_.clear(); keys_.clear(); order_.clear(); offsets_.clear();
    stats_ = Stats();
    return 0;
  }

  int train(IndexThreads::Pointer threads, IndexHolder::Pointer holder) override {
    if (stock_) return stock_->train(std::move(threads), std::move(holder));
    if (state_ != INITED) return IndexError_Runtime;
    if (!holder || holder->count() == 0) return IndexError_InvalidArgument;
    if (!holder->is_matched(meta_)) return IndexError_Mismatch;
    // the training sample (StratifiedClusterTrainer::train, stratified_cluster_trainer.cc:156-182): every s-th row
    const size_t total = holder->count();
    size_t want = std::max<size_t>(sample_count_, (size_t)(sample_ratio_ * total));
    if (want == 0 || want > total) want = total;
    want = std::max<size_t>(want, std::min<size_t>(total, nlist_));
    const size_t es = meta_.element_size(), dim = meta_.dimension();
    const size_t d = cosine_ ? dim - 1 : dim;                 // the norm slot does not enter the distance
    std::vector<float> sample;
    sample.reserve(want * d);
    size_t seen = 0, taken = 0;
    double norm_sum = 0.0;
    for (auto it = holder->create_iterator(); it && it->is_valid() && taken < want; it->next(), ++seen) {
      if (seen * want / total < taken) continue;             // row `seen` is not the next of `want` evenly spaced ones
      const float *v = static_cast<const float *>(it->data());
      sample.insert(sample.end(), v, v + d);
      if (cosine_) norm_sum += v[d];
      ++taken;
    }
    if (taken < nlist_) return IndexError_InvalidArgument;
    (void)es;
    // initial centroids: evenly spaced rows of the sample (KmeansCluster::init_centroids draws a reservoir sample)
    std::vector<float> cent((size_t)nlist_ * d);
    for (uint32_t c = 0; c < nlist_; ++c)
      std::memcpy(&cent[(size_t)c * d], &sample[((size_t)c * taken / nlist_) * d], d * sizeof(float));
    std::vector<uint64_t> follows(nlist_);
    uint32_t iters = 0;
    double sse = 0.0;
    // Cosine on normalised rows: assignment by the largest inner product with the cluster mean, which is what the
    // reference's clusters do under the Cosine metric (1 - x.c, the mean is not re-normalised)
    int rc = zvb200_kmeans(device_, cosine_ ? ZVB200_METRIC_IP : ZVB200_METRIC_L2SQ, sample.data(), taken, (uint32_t)d,
                           cent.data(), nlist_, max_iterations_, 1.1920929e-07, nullptr, follows.data(), &sse, &iters);
    if (rc != 0) { LOG_ERROR("GpuIVFBuilder: zvb200_kmeans failed: %s", zvb200_last_error()); return rc; }
    // clusters nobody follows sit at the tail (PurgeCentroids, kmeans_cluster.cc:363-383) and are cut, as
    // purge_empty does
    uint32_t live = nlist_;
    while (live > 1 && follows[live - 1] == 0) --live;
    IndexCluster::CentroidList list;
    std::vector<float> row(dim);
    for (uint32_t c = 0; c < live; ++c) {
      std::memcpy(row.data(), &cent[(size_t)c * d], d * sizeof(float));
      if (cosine_) row[d] = (float)(norm_sum / (double)taken);  // the slot of a mean row: the mean norm; never read by the metric
      list.emplace_back(row.data(), dim * sizeof(float));
    }
    stats_.set_trained_count(taken);
    return this->adopt(list);
  }

  int train(const IndexTrainer::Pointer &trainer) override {
    if (stock_) return stock_->train(trainer);
    if (state_ != INITED) return IndexError_Runtime;
    if (!trainer) return IndexError_InvalidArgument;
    IndexCluster::CentroidList list;
    int ret = IndexCluster::Deserialize(trainer->meta(), trainer->indexes(), &list);
    if (ret != 0) return ret;
    const IndexMeta &m = trainer->meta();
    if (m.data_type() != inner_meta_.data_type() || m.metric_name() != inner_meta_.metric_name() ||
        m.element_size() != inner_meta_.element_size())
      return IndexError_Mismatch;
    return this->adopt(list);
  }

  int build(IndexThreads::Pointer threads, IndexHolder::Pointer holder) override {
    if (stock_) return stock_->build(std::move(threads), std::move(holder));
    if (state_ != TRAINED) return IndexError_Runtime;
    if (!holder || holder->count() == 0) return IndexError_InvalidArgument;
    if (!holder->is_matched(meta_)) return IndexError_Mismatch;
    const size_t es = meta_.element_size();
    rows_.clear(); keys_.clear();
    rows_.reserve(holder->count() * es);
    keys_.reserve(holder->count());
    for (auto it = holder->create_iterator(); it && it->is_valid(); it->next()) {
      const char *v = static_cast<const char *>(it->data());
      rows_.insert(rows_.end(), v, v + es);
      keys_.push_back(it->key());
    }
    const uint64_t n = keys_.size();
    const uint32_t nl = (uint32_t)(centroids_.size() / es);
    std::vector<uint32_t> labels(n);
    int rc = zvb200_ivf_assign(device_, ZVB200_DT_FP32, cosine_ ? ZVB200_METRIC_COSINE : ZVB200_METRIC_L2SQ,
                               meta_.dimension(), centroids_.data(), nl, rows_.data(), n, labels.data());
    if (rc != 0) { LOG_ERROR("GpuIVFBuilder: zvb200_ivf_assign failed: %s", zvb200_last_error()); return rc; }
    order_.assign(n, 0);
    offsets_.assign((size_t)nl + 1, 0);
    rc = zvb200_ivf_pack(device_, labels.data(), n, nl, order_.data(), offsets_.data());
    if (rc != 0) { LOG_ERROR("GpuIVFBuilder: zvb200_ivf_pack failed: %s", zvb200_last_error()); return rc; }
    stats_.set_built_count(n);
    state_ = BUILT;
    return 0;
  }

  int dump(const IndexDumper::Pointer &dumper) override {
    if (stock_) return stock_->dump(dumper);
    if (state_ != BUILT) return IndexError_Runtime;
    IndexMeta list_meta = inner_meta_;
    int ret = this->choose_major_order(&list_meta);
    if (ret != 0) return ret;
    const uint32_t nl = (uint32_t)centroid_index_->centroids_count();
    const size_t es = meta_.element_size();
    auto ivf_dumper = std::make_shared<IVFDumper>(list_meta, dumper, nl, block_vector_count_);
    for (uint32_t l = 0; l < nl; ++l)
      for (uint64_t p = offsets_[l]; p < offsets_[l + 1]; ++p) {
        const uint32_t id = order_[p];
        ret = ivf_dumper->dump_inverted_vector(l, keys_[id], &rows_[(size_t)id * es]);
        if (ret != 0) return ret;
      }
    ret = ivf_dumper->dump_inverted_vector_finished();
    if (ret != 0) return ret;
    ret = ivf_dumper->dump_quantizer_params(std::vector<IndexConverter::Pointer>());
    if (ret != 0) return ret;
    ret = ivf_dumper->dump_centroid_index(centroid_index_->data(), centroid_index_->size());
    if (ret != 0) return ret;
    stats_.set_dumped_count(ivf_dumper->dumped_count());
    // the searcher's default scan ratio, fitted as IVFBuilder::dump does (ivf_builder.cc:416-425)
    float scan_ratio = -0.004f * std::log((float)keys_.size()) + 0.0751f;
    scan_ratio = std::max(scan_ratio, 0.0001f);
    ailego::Params sp;
    sp.set(PARAM_IVF_SEARCHER_SCAN_RATIO, scan_ratio);
    meta_.set_searcher("IVFSearcher", 0, std::move(sp));
    meta_.set_builder("GpuIVFBuilder", 0, ailego::Params(params_));
    ret = IndexHelper::SerializeToDumper(meta_, dumper.get());
    if (ret != 0) return ret;
    stats_.set_discarded_count(stats_.built_count() - stats_.dumped_count());
    return 0;
  }

  const Stats &stats(void) const override { return stock_ ? stock_->stats() : stats_; }

 private:
  enum State { INIT = 0, INITED = 1, TRAINED = 2, BUILT = 3 };

  //! what the device path builds itself
  bool accelerated(const IndexMeta &meta, const ailego::Params &params) const {
    if (meta.data_type() != IndexMeta::DataType::DT_FP32) return false;
    const std::string &m = meta.metric_name();
    if (m != "SquaredEuclidean" && m != "Cosine") return false;   // InnerProduct trains through the Mips converter
    if (m == "Cosine" && meta.dimension() < 2) return false;
    const std::string cn = params.get_as_string(PARAM_IVF_BUILDER_CENTROID_COUNT);
    if (cn.empty() || cn.find_first_not_of("0123456789") != std::string::npos) return false;   // "a*b": two levels
    for (const std::string *key : {&PARAM_IVF_BUILDER_CONVERTER_CLASS, &PARAM_IVF_BUILDER_QUANTIZER_CLASS,
                                   &PARAM_IVF_BUILDER_OPTIMIZER_CLASS, &PARAM_IVF_BUILDER_OPTIMIZER_QUANTIZER_CLASS})
      if (!params.get_as_string(*key).empty()) return false;
    if (params.get_as_bool(PARAM_IVF_BUILDER_STORE_ORIGINAL_FEATURES)) return false;
    if (params.get_as_bool(PARAM_IVF_BUILDER_QUANTIZE_BY_CENTROID)) return false;
    return zvb200_device_count() > 0;
  }

  //! the tail of IVFBuilder::train(trainer) (ivf_builder.cc:305-312): the reference's own centroid index
  int adopt(const IndexCluster::CentroidList &list) {
    if (list.empty()) return IndexError_InvalidArgument;
    const size_t es = meta_.element_size();
    centroids_.clear();
    for (const auto &c : list) {
      if (c.size() != es) return IndexError_Mismatch;
      const char *p = static_cast<const char *>(c.feature());
      centroids_.insert(centroids_.end(), p, p + es);
    }
    centroid_index_ = std::make_shared<IVFCentroidIndex>();
    int ret = centroid_index_->init(inner_meta_, params_);
    if (ret != 0) return ret;
    ret = centroid_index_->build(list);
    if (ret != 0) return ret;
    state_ = TRAINED;
    return 0;
  }

  //! IVFBuilder::CheckAndUpdateMajorOrder (ivf_builder.cc:440-480)
  int choose_major_order(IndexMeta *meta) const {
    auto metric = IndexFactory::CreateMetric(meta->metric_name());
    if (!metric) return IndexError_InvalidArgument;
    int ret = metric->init(*meta, meta->metric_params());
    if (ret != 0) return ret;
    bool column_ok = true;
    for (size_t m = 32; m != 0 && column_ok; m /= 2)
      for (size_t n = m; n != 0; n /= 2)
        if (metric->distance_matrix(m, n) == nullptr) { column_ok = false; break; }
    column_ok &= meta->element_size() % IndexMeta::AlignSizeof(meta->data_type()) == 0;
    if (meta->major_order() == IndexMeta::MO_UNDEFINED)
      meta->set_major_order(column_ok && meta->dimension() <= 512 ? IndexMeta::MO_COLUMN : IndexMeta::MO_ROW);
    else if (!column_ok && meta->major_order() == IndexMeta::MO_COLUMN)
      return IndexError_Unsupported;
    if (block_vector_count_ * meta->element_size() % 32 != 0) return IndexError_InvalidArgument;
    return 0;
  }

  State state_{INIT};
  Stats stats_{};
  IndexMeta meta_{}, inner_meta_{};
  ailego::Params params_{};
  IndexBuilder::Pointer stock_{};
  IVFCentroidIndex::Pointer centroid_index_{};
  uint32_t nlist_{0}, sample_count_{0}, block_vector_count_{32}, max_iterations_{20};
  float sample_ratio_{0.f};
  int device_{0};
  bool cosine_{false};
  std::vector<char> centroids_{}, rows_{};
  std::vector<uint64_t> keys_{}, offsets_{};
  std::vector<uint32_t> order_{};
};

INDEX_FACTORY_REGISTER_BUILDER(GpuIVFBuilder);

}  // namespace core
}  // namespace zvec
