// b2serve -- the model-server binary that takes the place of the vLLM container the reference's
// llm-d deployment launches (llm-d-deploy.yaml:176-184) and that llm-d-test.yaml:32-43,61-73 talks to:
//   GET  /v1/models            body contains the served model id (the reference's only assertion)
//   POST /v1/completions       prompt: str | [str] | [int] | [[int]]
//   POST /v1/chat/completions  messages rendered through --chat-template (templates/*.yaml)
//   GET  /health, /metrics     K8s probes; Prometheus text with vllm:* names on the serving port
// One GPU-owner thread drives the engine's C ABI (continuous batching: admit -> chunked prefill ->
// one decode step over all running sequences); HTTP connection threads only queue requests and
// stream tokens back.  Flags mirror the `vllm serve` subset a chart passes.
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <fstream>
#include <iostream>
#include <memory>
#include <mutex>
#include <random>
#include <sstream>
#include <thread>

#include "../../../include/b2_engine.h"
#include "../json.hpp"
#include "http.hpp"
#include "jinja.hpp"
#include "metrics.hpp"
#include "tokenizer.hpp"

using b2json::Value;

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static int64_t unix_s() { return (int64_t)std::chrono::duration_cast<std::chrono::seconds>(std::chrono::system_clock::now().time_since_epoch()).count(); }

// ------------------------------------------------------------------------------------------
// options
// ------------------------------------------------------------------------------------------
struct Options {
  std::string model = "Qwen/Qwen3-0.6B";
  std::string served_name;
  std::string host = "0.0.0.0";
  int port = 8000;
  std::string chat_template_path;
  std::string tokenizer_path;        // tokenizer.json, or a directory holding one; default: the model directory
  std::string load_format = "auto";  // auto: safetensors if the model dir has one, else dummy
  int max_model_len = 4096;
  int block_size = 16;
  float gpu_mem_util = 0.9f;
  int max_num_seqs = 256;
  int max_batched_tokens = 8192;
  int num_blocks = 0;
  uint64_t seed = 0;
  int eos_token_id = 151645;
  int device = 0;
  int use_graphs = 1;
  bool prefix_caching = false;  // --enable-prefix-caching
};

static void usage() {
  fprintf(stderr,
          "usage: b2serve [MODEL] [--model M] [--served-model-name N] [--host H] [--port P] [--chat-template FILE]\n"
          "               [--max-model-len L] [--block-size 16] [--gpu-memory-utilization F] [--max-num-seqs N]\n"
          "               [--max-num-batched-tokens T] [--num-gpu-blocks-override N] [--seed S] [--load-format auto|dummy|safetensors]\n"
          "               [--eos-token-id ID] [--device D] [--enforce-eager] [--tokenizer FILE_OR_DIR] [--enable-prefix-caching]\n"
          "       b2serve --render-chat-template FILE < {\"messages\": ...}     (no GPU)\n"
          "       b2serve --tokenize TOKENIZER_JSON < {\"text\": ...} per line  (no GPU)\n");
}

static bool parse_args(int argc, char** argv, Options& o) {
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto val = [&](const char* name) -> const char* {
      if (i + 1 >= argc) { fprintf(stderr, "b2serve: %s needs a value\n", name); exit(2); }
      return argv[++i];
    };
    std::string key = a, inl;
    size_t eq = a.find('=');
    bool has_inl = false;
    if (a.rfind("--", 0) == 0 && eq != std::string::npos) { key = a.substr(0, eq); inl = a.substr(eq + 1); has_inl = true; }
    auto V = [&](const char* name) -> std::string { return has_inl ? inl : std::string(val(name)); };
    if (key == "--model") o.model = V("--model");
    else if (key == "--served-model-name") o.served_name = V("--served-model-name");
    else if (key == "--host") o.host = V("--host");
    else if (key == "--port") o.port = atoi(V("--port").c_str());
    else if (key == "--chat-template") o.chat_template_path = V("--chat-template");
    else if (key == "--max-model-len") o.max_model_len = atoi(V("--max-model-len").c_str());
    else if (key == "--block-size") o.block_size = atoi(V("--block-size").c_str());
    else if (key == "--gpu-memory-utilization") o.gpu_mem_util = (float)atof(V("--gpu-memory-utilization").c_str());
    else if (key == "--max-num-seqs") o.max_num_seqs = atoi(V("--max-num-seqs").c_str());
    else if (key == "--max-num-batched-tokens") o.max_batched_tokens = atoi(V("--max-num-batched-tokens").c_str());
    else if (key == "--num-gpu-blocks-override") o.num_blocks = atoi(V("--num-gpu-blocks-override").c_str());
    else if (key == "--seed") o.seed = strtoull(V("--seed").c_str(), nullptr, 10);
    else if (key == "--load-format") o.load_format = V("--load-format");
    else if (key == "--eos-token-id") o.eos_token_id = atoi(V("--eos-token-id").c_str());
    else if (key == "--device") o.device = atoi(V("--device").c_str());
    else if (key == "--enforce-eager") o.use_graphs = 0;
    else if (key == "--help" || key == "-h") { usage(); exit(0); }
    else if (key == "--tokenizer") o.tokenizer_path = V("--tokenizer");
    else if (key == "--enable-prefix-caching") o.prefix_caching = true;
    else if (key == "--no-enable-prefix-caching") o.prefix_caching = false;
    else if (key == "--dtype" || key == "--tensor-parallel-size" || key == "--trust-remote-code" || key == "--disable-log-requests" ||
             key == "--kv-cache-dtype" || key == "--download-dir") {
      // accepted for chart compatibility; this build is bf16, one GPU per replica
      if (!has_inl && key != "--trust-remote-code" && key != "--disable-log-requests" && i + 1 < argc) ++i;
    } else if (a.rfind("--", 0) == 0) { fprintf(stderr, "b2serve: unknown flag %s\n", a.c_str()); return false; }
    else if (a == "serve") continue;  // `vllm serve MODEL` style
    else o.model = a;
  }
  if (o.served_name.empty()) o.served_name = o.model;
  return true;
}

// ------------------------------------------------------------------------------------------
// model config + weights
// ------------------------------------------------------------------------------------------
static std::string read_file(const std::string& p) {
  std::ifstream f(p, std::ios::binary);
  if (!f) return std::string();
  std::stringstream ss;
  ss << f.rdbuf();
  return ss.str();
}
static bool is_dir(const std::string& p) { struct stat st; return stat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode); }
static bool is_file(const std::string& p) { struct stat st; return stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode); }

static void fill_config(const Options& o, b2_config& c) {
  memset(&c, 0, sizeof c);
  c.abi_version = B2_ABI_VERSION;
  // Qwen3-0.6B (BASELINE.json north_star) unless the model directory carries a config.json
  c.num_layers = 28; c.hidden = 1024; c.num_q_heads = 16; c.num_kv_heads = 8; c.head_dim = 128; c.intermediate = 3072;
  c.vocab = 151936; c.tie_embeddings = 1; c.rms_eps = 1e-6f; c.rope_theta = 1e6f;
  if (is_dir(o.model) && is_file(o.model + "/config.json")) {
    Value j = b2json::parse(read_file(o.model + "/config.json"));
    auto geti = [&](const char* k, int d) { const Value* v = j.find(k); return v && v->is_number() ? (int)v->as_int() : d; };
    auto getf = [&](const char* k, double d) { const Value* v = j.find(k); return v && v->is_number() ? v->as_double() : d; };
    c.num_layers = geti("num_hidden_layers", c.num_layers);
    c.hidden = geti("hidden_size", c.hidden);
    c.num_q_heads = geti("num_attention_heads", c.num_q_heads);
    c.num_kv_heads = geti("num_key_value_heads", c.num_kv_heads);
    c.head_dim = geti("head_dim", c.hidden / std::max(1, c.num_q_heads));
    c.intermediate = geti("intermediate_size", c.intermediate);
    c.vocab = geti("vocab_size", c.vocab);
    const Value* tie = j.find("tie_word_embeddings");
    if (tie && tie->is_bool()) c.tie_embeddings = tie->b ? 1 : 0;
    c.rms_eps = (float)getf("rms_norm_eps", c.rms_eps);
    double theta = getf("rope_theta", c.rope_theta);
    const Value* rp = j.find("rope_parameters");
    if (rp && rp->find("rope_theta")) theta = rp->find("rope_theta")->as_double(theta);
    c.rope_theta = (float)theta;
    // model family: "model_type" of the HF config.json (qwen3 / phi3)
    const Value* mt = j.find("model_type");
    if (mt && mt->is_string() && mt->as_string() == "phi3") {
      c.arch = B2_ARCH_PHI3;
      c.rotary_dim = (int)(getf("partial_rotary_factor", 1.0) * c.head_dim + 0.5);
      c.sliding_window = geti("sliding_window", 0);  // null in the config parses as "not a number" -> 0
    } else if (mt && mt->is_string() && mt->as_string() == "opt") {
      // HF OPTConfig names: ffn_dim, no kv-head count (MHA), LayerNorm eps is torch's default
      c.arch = B2_ARCH_OPT;
      c.num_kv_heads = c.num_q_heads;
      c.head_dim = c.hidden / std::max(1, c.num_q_heads);
      c.intermediate = geti("ffn_dim", c.intermediate);
      c.max_positions = geti("max_position_embeddings", 2048);
      c.rms_eps = 1e-5f;
      c.tie_embeddings = 1;
    }
  }
  c.max_model_len = o.max_model_len;
  c.block_size = o.block_size;
  c.num_blocks = o.num_blocks;
  c.gpu_mem_fraction = o.gpu_mem_util;
  c.max_seqs = o.max_num_seqs;
  c.max_batch_tokens = std::max(o.max_batched_tokens, 16);
  c.use_cuda_graphs = o.use_graphs;
  c.rope_inv_freq = nullptr;
}

static uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return (uint16_t)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);
}

// HF default init statistics: N(0, 0.02) for Linear/Embedding, ones for norm gains.
static int load_dummy_weights(b2_engine* e, const b2_config& c, uint64_t seed) {
  struct T { std::string name; size_t n; bool norm; };
  std::vector<T> ts;
  const size_t q = (size_t)c.num_q_heads * c.head_dim, kv = (size_t)c.num_kv_heads * c.head_dim, H = c.hidden, I = c.intermediate;
  ts.push_back({"model.embed_tokens.weight", (size_t)c.vocab * H, false});
  for (int l = 0; l < c.num_layers; ++l) {
    std::string p = "model.layers." + std::to_string(l) + ".";
    ts.push_back({p + "input_layernorm.weight", H, true});
    ts.push_back({p + "self_attn.q_proj.weight", q * H, false});
    ts.push_back({p + "self_attn.k_proj.weight", kv * H, false});
    ts.push_back({p + "self_attn.v_proj.weight", kv * H, false});
    ts.push_back({p + "self_attn.o_proj.weight", H * q, false});
    ts.push_back({p + "self_attn.q_norm.weight", (size_t)c.head_dim, true});
    ts.push_back({p + "self_attn.k_norm.weight", (size_t)c.head_dim, true});
    ts.push_back({p + "post_attention_layernorm.weight", H, true});
    ts.push_back({p + "mlp.gate_proj.weight", I * H, false});
    ts.push_back({p + "mlp.up_proj.weight", I * H, false});
    ts.push_back({p + "mlp.down_proj.weight", H * I, false});
  }
  ts.push_back({"model.norm.weight", H, true});
  if (!c.tie_embeddings) ts.push_back({"lm_head.weight", (size_t)c.vocab * H, false});
  std::vector<uint16_t> buf;
  const unsigned nthreads = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
  for (size_t ti = 0; ti < ts.size(); ++ti) {
    const T& t = ts[ti];
    buf.resize(t.n);
    if (t.norm) {
      std::fill(buf.begin(), buf.end(), f2bf(1.0f));
    } else {
      std::vector<std::thread> th;
      for (unsigned k = 0; k < nthreads; ++k)
        th.emplace_back([&, k] {
          std::mt19937_64 rng(seed * 1000003ull + ti * 131ull + k);
          std::normal_distribution<float> nd(0.f, 0.02f);
          const size_t a = t.n * k / nthreads, b = t.n * (k + 1) / nthreads;
          for (size_t i = a; i < b; ++i) buf[i] = f2bf(nd(rng));
        });
      for (auto& x : th) x.join();
    }
    int rc = b2_engine_load_tensor(e, t.name.c_str(), buf.data(), t.n * 2);
    if (rc != 0) return rc;
  }
  return b2_engine_finalize(e);
}

// ------------------------------------------------------------------------------------------
// requests + scheduler
// ------------------------------------------------------------------------------------------
struct Req {
  std::vector<int32_t> prompt;
  int max_tokens = 16;
  b2_sampling samp{1.0f, 0, 1.0f, 0};
  bool ignore_eos = false;
  bool streaming = false;  // a handler is waiting for every token (SSE); otherwise only for the end
  std::vector<std::string> stops;
  // produced
  std::mutex mu;
  std::condition_variable cv;
  std::vector<int32_t> out_ids;
  std::string text;          // detokenised output so far (stop string already cut off)
  bool done = false;
  std::string finish_reason;  // "stop" | "length"
  std::string error;
  // scheduler-private
  int seq = -1;
  int prefilled = 0;
  int32_t next_token = 0;
  b2tok::Detok detok;
  double t_arrival = 0, t_first = 0, t_last = 0;
  std::atomic<bool> cancelled{false};
};
using ReqP = std::shared_ptr<Req>;

class Scheduler {
 public:
  Scheduler(b2_engine* e, const b2_config& c, const Options& o, b2metrics::Registry* m) : e_(e), cfg_(c), opt_(o), m_(m) {}
  void start() { th_ = std::thread([this] { loop(); }); }
  void stop() {
    { std::lock_guard<std::mutex> l(mu_); stop_ = true; }
    cv_.notify_all();
    if (th_.joinable()) th_.join();
  }
  void submit(const ReqP& r) {
    r->t_arrival = now_s();
    { std::lock_guard<std::mutex> l(mu_); waiting_.push_back(r); m_->waiting = (int64_t)waiting_.size(); }
    cv_.notify_all();
  }

 private:
  b2_engine* e_;
  b2_config cfg_;
  Options opt_;
  b2metrics::Registry* m_;
  std::thread th_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<ReqP> waiting_;
  std::vector<ReqP> prefilling_, running_;
  bool stop_ = false;
  bool logits_on_ = false;

  void fail(const ReqP& r, const std::string& msg) {
    std::lock_guard<std::mutex> l(r->mu);
    r->error = msg;
    r->done = true;
    r->cv.notify_all();
  }
  void release(const ReqP& r) {
    if (r->seq >= 0) { b2_seq_free(e_, r->seq); r->seq = -1; }
  }
  void finish(const ReqP& r, const char* reason) {
    release(r);
    const double t = now_s();
    m_->e2e.observe(t - r->t_arrival);
    if (std::string(reason) == "stop") m_->success_stop++; else m_->success_length++;
    std::lock_guard<std::mutex> l(r->mu);
    r->text += r->detok.finish();
    r->finish_reason = reason;
    r->done = true;
    r->cv.notify_all();
  }
  // returns true when the request finished with this token
  bool deliver(const ReqP& r, int32_t tok) {
    const double t = now_s();
    if (r->out_ids.empty()) { r->t_first = t; m_->ttft.observe(t - r->t_arrival); }
    else m_->itl.observe(t - r->t_last);
    r->t_last = t;
    m_->generation_tokens++;
    bool eos = (!r->ignore_eos && tok == opt_.eos_token_id);
    bool stop_hit = false;
    {
      std::lock_guard<std::mutex> l(r->mu);
      r->out_ids.push_back(tok);
      if (!eos) {
        const size_t before = r->text.size();
        r->text += r->detok.push(tok);
        for (auto& s : r->stops) {
          if (s.empty()) continue;
          const size_t from = before > s.size() ? before - s.size() : 0;
          const size_t pos = r->text.find(s, from);
          if (pos != std::string::npos) { r->text.erase(pos); stop_hit = true; break; }
        }
      }
      // a non-streaming handler only cares about `done`: waking 64 of them per decode step costs the scheduler
      // thread more host time than the step's own bookkeeping
      if (r->streaming) r->cv.notify_all();
    }
    r->next_token = tok;
    if (eos || stop_hit) { finish(r, "stop"); return true; }
    if ((int)r->out_ids.size() >= r->max_tokens) { finish(r, "length"); return true; }
    if (r->cancelled) { release(r); m_->success_abort++; std::lock_guard<std::mutex> l(r->mu); r->done = true; r->finish_reason = "abort"; r->cv.notify_all(); return true; }
    return false;
  }
  bool needs_logits() const {
    for (auto& r : running_) if (r->samp.temperature > 0.f) return true;
    for (auto& r : prefilling_) if (r->samp.temperature > 0.f) return true;
    return false;
  }
  void update_gauges() {
    int64_t nb = 1, fb = 0;
    b2_engine_info(e_, &nb, &fb, nullptr);
    m_->kv_total_blocks = std::max<int64_t>(1, nb - 1);
    m_->kv_used_blocks = std::max<int64_t>(0, nb - 1 - fb);
    int64_t q = 0, h = 0;
    if (b2_prefix_cache_stats(e_, &q, &h, nullptr) == 0) { m_->prefix_queries = q; m_->prefix_hits = h; }
    m_->running = (int64_t)(running_.size() + prefilling_.size());
  }

  void loop() {
    b2_sampling greedy{0.f, 0, 1.f, 0};
    (void)greedy;
    while (true) {
      // ---- admit ----
      {
        std::unique_lock<std::mutex> l(mu_);
        if (running_.empty() && prefilling_.empty() && waiting_.empty()) {
          update_gauges();
          cv_.wait(l, [this] { return stop_ || !waiting_.empty(); });
        }
        if (stop_) break;
        while (!waiting_.empty() && (int)(running_.size() + prefilling_.size()) < cfg_.max_seqs) {
          ReqP r = waiting_.front();
          const int need = std::min(cfg_.max_model_len, (int)r->prompt.size() + r->max_tokens);
          int32_t sid = -1;
          if (b2_seq_create(e_, need, &sid) != 0) {
            // Pool full: retry after sequences finish -- unless nothing is running, in which case no page will ever
            // come back and this request (it needs more pages than the pool holds) would block the queue for good.
            if (running_.empty() && prefilling_.empty()) {
              waiting_.pop_front();
              fail(r, std::string("request needs ") + std::to_string(need) + " tokens of KV cache, more than the pool holds: " +
                          b2_last_error());
              continue;
            }
            break;
          }
          waiting_.pop_front();
          r->seq = sid;
          prefilling_.push_back(r);
        }
        m_->waiting = (int64_t)waiting_.size();
      }
      const bool want_logits = needs_logits();
      if (want_logits != logits_on_) { b2_set_debug(e_, want_logits ? 2 : 0); logits_on_ = want_logits; }

      // ---- chunked prefill: pack admitted prompts into ONE forward of <= max_batch_tokens tokens per loop turn, so the
      //      running streams get a decode step between the chunks of a long prompt ----
      if (!prefilling_.empty()) {
        std::vector<int32_t> ids, offs{0}, sids;
        std::vector<ReqP> batch;
        std::vector<int> take;
        int budget = cfg_.max_batch_tokens;
        for (auto& r : prefilling_) {
          if (budget <= 0 || (int)batch.size() >= cfg_.max_seqs) break;
          const int left = (int)r->prompt.size() - r->prefilled;
          const int n = std::min(left, budget);
          ids.insert(ids.end(), r->prompt.begin() + r->prefilled, r->prompt.begin() + r->prefilled + n);
          offs.push_back((int32_t)ids.size());
          sids.push_back(r->seq);
          batch.push_back(r);
          take.push_back(n);
          budget -= n;
        }
        std::vector<int32_t> out(batch.size());
        int rc = b2_prefill(e_, ids.data(), offs.data(), (int32_t)batch.size(), sids.data(), out.data());
        if (rc != 0) {
          const std::string msg = b2_last_error();
          for (auto& r : batch) { release(r); fail(r, msg); }
          prefilling_.erase(prefilling_.begin(), prefilling_.begin() + batch.size());
          continue;
        }
        // sampled first tokens for non-greedy requests whose prompt is now complete
        std::vector<int32_t> rows;
        std::vector<b2_sampling> sp;
        for (size_t i = 0; i < batch.size(); ++i)
          if (batch[i]->prefilled + take[i] == (int)batch[i]->prompt.size() && batch[i]->samp.temperature > 0.f) { rows.push_back((int32_t)i); sp.push_back(batch[i]->samp); }
        std::vector<int32_t> sampled(rows.size());
        if (!rows.empty() && b2_sample_rows(e_, (int32_t)rows.size(), rows.data(), sp.data(), sampled.data()) == 0)
          for (size_t k = 0; k < rows.size(); ++k) out[rows[k]] = sampled[k];
        std::vector<ReqP> still;
        for (size_t i = 0; i < batch.size(); ++i) {
          ReqP& r = batch[i];
          r->prefilled += take[i];
          m_->prompt_tokens += take[i];
          if (r->prefilled < (int)r->prompt.size()) { still.push_back(r); continue; }
          if (!deliver(r, out[i])) running_.push_back(r);
        }
        prefilling_.erase(prefilling_.begin(), prefilling_.begin() + batch.size());
        prefilling_.insert(prefilling_.begin(), still.begin(), still.end());
      }

      // ---- one decode step over every running sequence ----
      if (!running_.empty()) {
        const int n = (int)running_.size();
        std::vector<int32_t> sids(n), in(n), out(n);
        for (int i = 0; i < n; ++i) { sids[i] = running_[i]->seq; in[i] = running_[i]->next_token; }
        int rc = b2_decode_step(e_, n, sids.data(), in.data(), out.data());
        if (rc != 0) {
          const std::string msg = b2_last_error();
          for (auto& r : running_) { release(r); fail(r, msg); }
          running_.clear();
          continue;
        }
        std::vector<int32_t> rows;
        std::vector<b2_sampling> sp;
        for (int i = 0; i < n; ++i)
          if (running_[i]->samp.temperature > 0.f) { rows.push_back(i); sp.push_back(running_[i]->samp); sp.back().seed += running_[i]->out_ids.size(); }
        std::vector<int32_t> sampled(rows.size());
        if (!rows.empty() && b2_sample_rows(e_, (int32_t)rows.size(), rows.data(), sp.data(), sampled.data()) == 0)
          for (size_t k = 0; k < rows.size(); ++k) out[rows[k]] = sampled[k];
        std::vector<ReqP> keep;
        for (int i = 0; i < n; ++i)
          if (!deliver(running_[i], out[i])) keep.push_back(running_[i]);
        running_.swap(keep);
      }
      update_gauges();
    }
    for (auto& r : running_) { release(r); fail(r, "server shutting down"); }
    for (auto& r : prefilling_) { release(r); fail(r, "server shutting down"); }
    for (auto& r : waiting_) fail(r, "server shutting down");
  }
};

// ------------------------------------------------------------------------------------------
// OpenAI-compatible API
// ------------------------------------------------------------------------------------------
struct App {
  Options opt;
  b2_config cfg;
  b2_engine* engine = nullptr;
  std::unique_ptr<Scheduler> sched;
  b2metrics::Registry metrics;
  std::unique_ptr<b2jinja::Template> chat_template;
  std::atomic<uint64_t> next_id{1};
  int64_t created = unix_s();
};

static const char* kDefaultChatTemplate =
    // ChatML, as Qwen's own template renders plain messages; no whitespace control: a `{%-` would eat the
    // newline that belongs to the prompt
    "{% for message in messages %}<|im_start|>{{ message['role'] }}\n{{ message['content'] }}<|im_end|>\n{% endfor %}"
    "{% if add_generation_prompt %}<|im_start|>assistant\n{% endif %}";

static std::string err_json(const std::string& msg, const std::string& type, int code) {
  Value e = Value::object();
  Value inner = Value::object();
  inner.set("message", msg);
  inner.set("type", type);
  inner.set("param", Value());
  inner.set("code", code);
  e.set("error", inner);
  return b2json::dump(e);
}
static void send_error(b2http::Conn& c, int status, const std::string& msg, const std::string& type = "invalid_request_error") {
  c.respond(status, "application/json", err_json(msg, type, status));
}

struct Parsed {
  std::vector<std::vector<int32_t>> prompts;
  int n = 1;
  int max_tokens = 16;
  bool max_tokens_given = false;
  b2_sampling samp{1.0f, 0, 1.0f, 0};
  bool stream = false, ignore_eos = false;
  std::vector<std::string> stops;
};

static bool parse_common(const Value& j, Parsed& p, std::string& err) {
  if (const Value* v = j.find("max_tokens")) if (v->is_number()) { p.max_tokens = (int)v->as_int(); p.max_tokens_given = true; }
  if (const Value* v = j.find("max_completion_tokens")) if (v->is_number()) { p.max_tokens = (int)v->as_int(); p.max_tokens_given = true; }
  if (const Value* v = j.find("temperature")) if (v->is_number()) p.samp.temperature = (float)v->as_double();
  if (const Value* v = j.find("top_p")) if (v->is_number()) p.samp.top_p = (float)v->as_double();
  if (const Value* v = j.find("top_k")) if (v->is_number()) p.samp.top_k = (int)v->as_int();
  if (const Value* v = j.find("seed")) if (v->is_number()) p.samp.seed = (uint64_t)v->as_int();
  if (const Value* v = j.find("n")) if (v->is_number()) p.n = (int)v->as_int();
  if (const Value* v = j.find("stream")) p.stream = v->as_bool(false);
  if (const Value* v = j.find("ignore_eos")) p.ignore_eos = v->as_bool(false);
  if (const Value* v = j.find("stop")) {
    if (v->is_string()) p.stops.push_back(v->str);
    else if (v->is_array()) for (size_t i = 0; i < v->size(); ++i) if (v->at(i).is_string()) p.stops.push_back(v->at(i).str);
  }
  if (p.samp.top_k < 0) p.samp.top_k = 0;
  if (p.max_tokens < 1) { err = "max_tokens must be at least 1"; return false; }
  if (p.n < 1 || p.n > 64) { err = "n must be between 1 and 64"; return false; }
  if (p.samp.temperature < 0.f) { err = "temperature must be non-negative"; return false; }
  if (p.samp.top_p <= 0.f || p.samp.top_p > 1.f) { err = "top_p must be in (0, 1]"; return false; }
  return true;
}

static bool model_ok(App& app, const Value& j, b2http::Conn& c) {
  const Value* m = j.find("model");
  if (m && m->is_string() && m->str != app.opt.served_name && m->str != app.opt.model) {
    c.respond(404, "application/json", err_json("The model `" + m->str + "` does not exist.", "NotFoundError", 404));
    return false;
  }
  return true;
}

static std::vector<ReqP> make_requests(App& app, const Parsed& p, std::string& err) {
  std::vector<ReqP> rs;
  for (auto& ids : p.prompts) {
    if (ids.empty()) { err = "prompt must not be empty"; return {}; }
    for (int32_t t : ids) if (t < 0 || t >= app.cfg.vocab) { err = "token id out of vocabulary range"; return {}; }
    if ((int)ids.size() >= app.cfg.max_model_len) { err = "prompt (" + std::to_string(ids.size()) + " tokens) exceeds the maximum model length of " + std::to_string(app.cfg.max_model_len); return {}; }
    for (int k = 0; k < p.n; ++k) {
      auto r = std::make_shared<Req>();
      r->prompt = ids;
      r->max_tokens = std::min(p.max_tokens, app.cfg.max_model_len - (int)ids.size());
      r->samp = p.samp;
      r->samp.seed = p.samp.seed * 0x9E3779B97F4A7C15ull + (uint64_t)k + (p.samp.seed == 0 ? app.next_id.load() : 0);
      r->ignore_eos = p.ignore_eos;
      r->stops = p.stops;
      rs.push_back(r);
    }
  }
  return rs;
}

static Value usage_json(const std::vector<ReqP>& rs, size_t prompts_each) {
  int64_t pt = 0, ct = 0;
  for (size_t i = 0; i < rs.size(); ++i) {
    if (i % prompts_each == 0) pt += (int64_t)rs[i]->prompt.size();
    ct += (int64_t)rs[i]->out_ids.size();
  }
  Value u = Value::object();
  u.set("prompt_tokens", pt);
  u.set("total_tokens", pt + ct);
  u.set("completion_tokens", ct);
  return u;
}

static void run_requests(App& app, std::vector<ReqP>& rs, bool stream, bool chat, const std::string& id, b2http::Conn& c, int n_per_prompt) {
  for (auto& r : rs) r->streaming = stream;
  for (auto& r : rs) app.sched->submit(r);
  const char* object = chat ? (stream ? "chat.completion.chunk" : "chat.completion") : "text_completion";
  auto base = [&]() {
    Value v = Value::object();
    v.set("id", id);
    v.set("object", object);
    v.set("created", unix_s());
    v.set("model", app.opt.served_name);
    return v;
  };
  if (!stream) {
    for (auto& r : rs) {
      std::unique_lock<std::mutex> l(r->mu);
      r->cv.wait(l, [&] { return r->done; });
    }
    for (auto& r : rs) if (!r->error.empty()) { send_error(c, 500, r->error, "internal_error"); return; }
    Value resp = base();
    Value choices = Value::array();
    for (size_t i = 0; i < rs.size(); ++i) {
      Value ch = Value::object();
      ch.set("index", (int64_t)i);
      if (chat) {
        Value msg = Value::object();
        msg.set("role", "assistant");
        msg.set("content", rs[i]->text);
        ch.set("message", msg);
      } else {
        ch.set("text", rs[i]->text);
      }
      ch.set("logprobs", Value());
      ch.set("finish_reason", rs[i]->finish_reason);
      choices.push(ch);
    }
    resp.set("choices", choices);
    resp.set("usage", usage_json(rs, (size_t)n_per_prompt));
    c.respond(200, "application/json", b2json::dump(resp));
    return;
  }
  // ---- SSE ----
  c.begin_sse();
  std::vector<size_t> sent(rs.size(), 0);
  std::vector<bool> closed(rs.size(), false);
  if (chat)
    for (size_t i = 0; i < rs.size(); ++i) {
      Value v = base();
      Value ch = Value::object();
      ch.set("index", (int64_t)i);
      Value d = Value::object();
      d.set("role", "assistant");
      d.set("content", "");
      ch.set("delta", d);
      ch.set("logprobs", Value());
      ch.set("finish_reason", Value());
      Value arr = Value::array();
      arr.push(ch);
      v.set("choices", arr);
      c.sse_event(b2json::dump(v));
    }
  size_t open = rs.size();
  while (open > 0 && c.ok()) {
    for (size_t i = 0; i < rs.size(); ++i) {
      if (closed[i]) continue;
      std::string delta, fin, error;
      bool done;
      {
        std::unique_lock<std::mutex> l(rs[i]->mu);
        if (rs.size() == 1) rs[i]->cv.wait(l, [&] { return rs[i]->done || rs[i]->text.size() > sent[i]; });
        else rs[i]->cv.wait_for(l, std::chrono::milliseconds(2), [&] { return rs[i]->done || rs[i]->text.size() > sent[i]; });
        done = rs[i]->done;
        // hold back a tail that could still turn into a stop string
        size_t upto = rs[i]->text.size();
        if (!done) { size_t hold = 0; for (auto& s : rs[i]->stops) hold = std::max(hold, s.size()); upto = upto > hold ? upto - hold : 0; }
        if (upto > sent[i]) { delta = rs[i]->text.substr(sent[i], upto - sent[i]); sent[i] = upto; }
        fin = rs[i]->finish_reason;
        error = rs[i]->error;
      }
      if (!delta.empty() || done) {
        Value v = base();
        Value ch = Value::object();
        ch.set("index", (int64_t)i);
        if (chat) { Value d = Value::object(); if (!delta.empty()) d.set("content", delta); ch.set("delta", d); }
        else ch.set("text", delta);
        ch.set("logprobs", Value());
        ch.set("finish_reason", done ? (error.empty() ? Value(fin) : Value("error")) : Value());
        Value arr = Value::array();
        arr.push(ch);
        v.set("choices", arr);
        c.sse_event(b2json::dump(v));
      }
      if (done) { closed[i] = true; --open; }
    }
  }
  if (!c.ok()) for (auto& r : rs) r->cancelled = true;  // client went away
  c.sse_event("[DONE]");
  c.end_sse();
}

static void handle(App& app, const b2http::Request& req, b2http::Conn& c) {
  try {
    if (req.method == "GET" && (req.path == "/health" || req.path == "/ping")) { c.respond(200, "text/plain", ""); return; }
    if (req.method == "GET" && req.path == "/metrics") { c.respond(200, "text/plain; version=0.0.4; charset=utf-8", app.metrics.render()); return; }
    if (req.method == "GET" && req.path == "/version") { c.respond(200, "application/json", "{\"version\":\"b2serve-0.1\"}"); return; }
    if (req.method == "GET" && req.path == "/v1/models") {
      Value m = Value::object();
      m.set("id", app.opt.served_name);
      m.set("object", "model");
      m.set("created", app.created);
      m.set("owned_by", "b2serve");
      m.set("root", app.opt.model);
      m.set("parent", Value());
      m.set("max_model_len", app.cfg.max_model_len);
      Value data = Value::array();
      data.push(m);
      Value resp = Value::object();
      resp.set("object", "list");
      resp.set("data", data);
      c.respond(200, "application/json", b2json::dump(resp));
      return;
    }
    const bool is_cmpl = req.path == "/v1/completions", is_chat = req.path == "/v1/chat/completions";
    if (is_cmpl || is_chat) {
      if (req.method != "POST") { send_error(c, 405, "method not allowed"); return; }
      Value j;
      try { j = b2json::parse(req.body); } catch (const std::exception& ex) { send_error(c, 400, std::string("invalid JSON body: ") + ex.what()); return; }
      if (!j.is_object()) { send_error(c, 400, "request body must be a JSON object"); return; }
      if (!model_ok(app, j, c)) return;
      Parsed p;
      std::string err;
      if (!parse_common(j, p, err)) { send_error(c, 400, err); return; }
      if (is_cmpl) {
        const Value* pr = j.find("prompt");
        if (!pr) { send_error(c, 400, "`prompt` is required"); return; }
        if (pr->is_string()) p.prompts.push_back(b2tok::encode(pr->str));
        else if (pr->is_array() && pr->size() > 0) {
          if (pr->at(0).is_number()) {
            std::vector<int32_t> ids;
            for (size_t i = 0; i < pr->size(); ++i) { if (!pr->at(i).is_number()) { send_error(c, 400, "mixed prompt array"); return; } ids.push_back((int32_t)pr->at(i).as_int()); }
            p.prompts.push_back(ids);
          } else {
            for (size_t i = 0; i < pr->size(); ++i) {
              const Value& e = pr->at(i);
              if (e.is_string()) p.prompts.push_back(b2tok::encode(e.str));
              else if (e.is_array()) { std::vector<int32_t> ids; for (size_t k = 0; k < e.size(); ++k) ids.push_back((int32_t)e.at(k).as_int()); p.prompts.push_back(ids); }
              else { send_error(c, 400, "prompt must be a string, a list of strings, a list of token ids or a list of token-id lists"); return; }
            }
          }
        } else { send_error(c, 400, "prompt must be a string, a list of strings, a list of token ids or a list of token-id lists"); return; }
      } else {
        const Value* msgs = j.find("messages");
        if (!msgs || !msgs->is_array() || msgs->size() == 0) { send_error(c, 400, "`messages` must be a non-empty list"); return; }
        for (size_t i = 0; i < msgs->size(); ++i)
          if (!msgs->at(i).is_object() || !msgs->at(i).find("role")) { send_error(c, 400, "every message needs a role and content"); return; }
        Value ctx = Value::object();
        ctx.set("messages", *msgs);
        const Value* agp = j.find("add_generation_prompt");
        ctx.set("add_generation_prompt", Value(agp ? agp->as_bool(true) : true));
        std::string text;
        try { text = app.chat_template->render(ctx); } catch (const std::exception& ex) { send_error(c, 400, std::string("chat template error: ") + ex.what()); return; }
        p.prompts.push_back(b2tok::encode(text));
        if (!p.max_tokens_given) p.max_tokens = std::max(1, app.cfg.max_model_len - (int)p.prompts[0].size());
      }
      std::vector<ReqP> rs = make_requests(app, p, err);
      if (rs.empty()) { send_error(c, 400, err); return; }
      const std::string id = std::string(is_chat ? "chatcmpl-" : "cmpl-") + std::to_string(app.next_id++);
      run_requests(app, rs, p.stream, is_chat, id, c, p.n);
      return;
    }
    send_error(c, 404, "not found: " + req.path, "NotFoundError");
  } catch (const std::exception& ex) {
    send_error(c, 500, ex.what(), "internal_error");
  }
}

// `b2serve --render-chat-template FILE` : reads {"messages": [...], "add_generation_prompt": bool} from
// stdin and prints the rendered prompt.  Needs no GPU; the CPU tests replay the golden renderings with it.
static int render_mode(const char* path) {
  std::string raw = read_file(path);
  if (raw.empty()) { fprintf(stderr, "b2serve: cannot read %s\n", path); return 1; }
  std::stringstream ss;
  ss << std::cin.rdbuf();
  try {
    b2jinja::Template t(b2jinja::extract_template(raw));
    Value ctx = b2json::parse(ss.str());
    std::string out = t.render(ctx);
    fwrite(out.data(), 1, out.size(), stdout);
  } catch (const std::exception& ex) {
    fprintf(stderr, "b2serve: %s\n", ex.what());
    return 1;
  }
  return 0;
}

// `b2serve --tokenize TOKENIZER_JSON` : reads one JSON object {"text": "..."} per line from stdin and prints
// {"ids": [...], "text": <ids decoded again>} per line.  Needs no GPU; the CPU tests compare it with the HF
// `tokenizers` library on the same tokenizer.json.
static int tokenize_mode(const char* path) {
  try {
    std::string raw = read_file(path);
    if (raw.empty()) throw std::runtime_error(std::string("cannot read ") + path);
    b2tok::global_bpe() = std::make_shared<b2tok::Bpe>(raw);
    std::string line;
    while (std::getline(std::cin, line)) {
      if (line.empty()) continue;
      Value in = b2json::parse(line);
      const Value* t = in.find("text");
      std::vector<int32_t> ids = b2tok::encode(t ? t->as_string() : std::string());
      b2tok::Detok d;
      std::string back;
      Value arr = Value::array();
      for (int32_t id : ids) { arr.push(Value((int64_t)id)); back += d.push(id); }
      back += d.finish();
      Value out = Value::object();
      out.set("ids", arr);
      out.set("text", back);
      std::string o = b2json::dump(out);
      fwrite(o.data(), 1, o.size(), stdout);
      fputc('\n', stdout);
    }
  } catch (const std::exception& ex) {
    fprintf(stderr, "b2serve: %s\n", ex.what());
    return 1;
  }
  return 0;
}

int main(int argc, char** argv) {
  if (argc == 3 && std::string(argv[1]) == "--render-chat-template") return render_mode(argv[2]);
  if (argc == 3 && std::string(argv[1]) == "--tokenize") return tokenize_mode(argv[2]);
  App app;
  if (!parse_args(argc, argv, app.opt)) { usage(); return 2; }
  fill_config(app.opt, app.cfg);
  // checkpoint files: model.safetensors, or the shards listed in model.safetensors.index.json (Phi-3-mini ships two)
  std::string st_path;
  std::vector<std::string> shards;
  if (app.opt.load_format != "dummy" && is_dir(app.opt.model) && is_file(app.opt.model + "/model.safetensors")) st_path = app.opt.model + "/model.safetensors";
  else if (app.opt.load_format != "dummy" && is_dir(app.opt.model) && is_file(app.opt.model + "/model.safetensors.index.json")) {
    try {
      Value idx = b2json::parse(read_file(app.opt.model + "/model.safetensors.index.json"));
      const Value* wm = idx.find("weight_map");
      if (wm && wm->is_object())
        for (auto& kv : *wm->obj) {
          const std::string f = app.opt.model + "/" + kv.second.as_string();
          if (std::find(shards.begin(), shards.end(), f) == shards.end()) shards.push_back(f);
        }
    } catch (const std::exception& ex) {
      fprintf(stderr, "b2serve: model.safetensors.index.json: %s\n", ex.what());
      return 1;
    }
  } else if (app.opt.load_format != "dummy" && is_file(app.opt.model)) st_path = app.opt.model;
  if (app.opt.load_format == "safetensors" && st_path.empty() && shards.empty()) { fprintf(stderr, "b2serve: no model.safetensors under %s\n", app.opt.model.c_str()); return 1; }
  const double t0 = now_s();
  int rc = b2_engine_create(&app.cfg, st_path.empty() ? nullptr : st_path.c_str(), app.opt.device, &app.engine);
  if (rc != 0) { fprintf(stderr, "b2serve: engine creation failed (%d): %s\n", rc, b2_last_error()); return 1; }
  if (!shards.empty()) {
    for (auto& f : shards) {
      rc = b2_engine_load_safetensors(app.engine, f.c_str());
      if (rc != 0) { fprintf(stderr, "b2serve: %s: %s\n", f.c_str(), b2_last_error()); return 1; }
    }
    rc = b2_engine_finalize(app.engine);
    if (rc != 0) { fprintf(stderr, "b2serve: %s\n", b2_last_error()); return 1; }
  } else if (st_path.empty()) {
    fprintf(stderr, "b2serve: no checkpoint found for %s; using random-init bf16 weights (seed %llu)\n", app.opt.model.c_str(), (unsigned long long)app.opt.seed);
    rc = load_dummy_weights(app.engine, app.cfg, app.opt.seed);
    if (rc != 0) { fprintf(stderr, "b2serve: weight init failed: %s\n", b2_last_error()); return 1; }
  }
  {  // tokenizer.json next to the weights (or --tokenizer): byte-level BPE; otherwise the byte fallback
    std::string tp = app.opt.tokenizer_path.empty() ? app.opt.model : app.opt.tokenizer_path;
    if (is_dir(tp)) tp += "/tokenizer.json";
    if (is_file(tp) && tp.size() > 5 && tp.compare(tp.size() - 5, 5, ".json") == 0) {
      try {
        b2tok::global_bpe() = std::make_shared<b2tok::Bpe>(read_file(tp));
        fprintf(stderr, "b2serve: tokenizer %s (%d tokens)\n", tp.c_str(), b2tok::global_bpe()->vocab_size());
      } catch (const std::exception& ex) {
        fprintf(stderr, "b2serve: %s: %s\n", tp.c_str(), ex.what());
        return 1;
      }
    } else if (!app.opt.tokenizer_path.empty()) {
      fprintf(stderr, "b2serve: no tokenizer.json at %s\n", app.opt.tokenizer_path.c_str());
      return 1;
    } else {
      fprintf(stderr, "b2serve: no tokenizer.json for %s; text prompts use the byte fallback (1 byte = 1 id)\n", app.opt.model.c_str());
    }
  }
  std::string tpl = kDefaultChatTemplate;
  if (!app.opt.chat_template_path.empty()) {
    std::string raw = read_file(app.opt.chat_template_path);
    if (raw.empty()) { fprintf(stderr, "b2serve: cannot read chat template %s\n", app.opt.chat_template_path.c_str()); return 1; }
    tpl = b2jinja::extract_template(raw);
  }
  try { app.chat_template.reset(new b2jinja::Template(tpl)); } catch (const std::exception& ex) { fprintf(stderr, "b2serve: chat template: %s\n", ex.what()); return 1; }
  app.metrics.model = app.opt.served_name;
  int64_t nb = 0, fb = 0, wb = 0;
  b2_engine_info(app.engine, &nb, &fb, &wb);
  if (app.opt.prefix_caching && b2_set_tuning(app.engine, B2_TUNE_PREFIX_CACHE, 1) != 0) {
    fprintf(stderr, "b2serve: --enable-prefix-caching: %s\n", b2_last_error());
    return 1;
  }
  app.sched.reset(new Scheduler(app.engine, app.cfg, app.opt, &app.metrics));
  app.sched->start();
  b2http::Server srv(app.opt.host, app.opt.port, [&app](const b2http::Request& r, b2http::Conn& c) { handle(app, r, c); });
  const int port = srv.listen_and_start();
  if (port < 0) { fprintf(stderr, "b2serve: cannot listen on %s:%d\n", app.opt.host.c_str(), app.opt.port); return 1; }
  fprintf(stderr, "b2serve: serving %s on %s:%d (%.1f s start-up, %.2f GB weights, %lld KV pages of %d tokens, max_model_len %d)\n",
          app.opt.served_name.c_str(), app.opt.host.c_str(), port, now_s() - t0, wb / 1e9, (long long)nb, app.cfg.block_size, app.cfg.max_model_len);
  printf("READY %d\n", port);
  fflush(stdout);
  // run until stdin closes or SIGTERM
  static std::atomic<bool> term{false};
  signal(SIGTERM, [](int) { term = true; });
  signal(SIGINT, [](int) { term = true; });
  while (!term) std::this_thread::sleep_for(std::chrono::milliseconds(100));
  srv.stop();
  app.sched->stop();
  b2_engine_free(app.engine);
  return 0;
}
