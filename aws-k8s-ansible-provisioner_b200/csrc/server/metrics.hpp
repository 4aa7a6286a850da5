// metrics.hpp -- Prometheus text exposition with the metric names the llm-d endpoint picker and the
// reference's OTel `vllm-metrics` scrape job read (otel-observability-setup.yaml:337-391; names as in
// vLLM v1/metrics/loggers.py:452-871): vllm:num_requests_running, vllm:num_requests_waiting,
// vllm:kv_cache_usage_perc (+ legacy vllm:gpu_cache_usage_perc), vllm:prompt_tokens_total,
// vllm:generation_tokens_total, vllm:request_success_total, and the TTFT / inter-token / e2e histograms.
#pragma once
#include <atomic>
#include <mutex>
#include <string>
#include <vector>

namespace b2metrics {

struct Histogram {
  std::vector<double> bounds;
  std::vector<uint64_t> counts;  // per bucket (non-cumulative), last = +Inf
  double sum = 0;
  uint64_t n = 0;
  std::mutex mu;
  explicit Histogram(std::vector<double> b) : bounds(std::move(b)), counts(bounds.size() + 1, 0) {}
  void observe(double v) {
    std::lock_guard<std::mutex> l(mu);
    size_t i = 0;
    while (i < bounds.size() && v > bounds[i]) ++i;
    counts[i]++;
    sum += v;
    n++;
  }
  void render(std::string& o, const std::string& name, const std::string& help, const std::string& labels) {
    std::lock_guard<std::mutex> l(mu);
    o += "# HELP " + name + " " + help + "\n# TYPE " + name + " histogram\n";
    uint64_t cum = 0;
    char b[128];
    for (size_t i = 0; i < bounds.size(); ++i) {
      cum += counts[i];
      snprintf(b, sizeof b, "%g", bounds[i]);
      o += name + "_bucket{" + labels + ",le=\"" + b + "\"} " + std::to_string(cum) + "\n";
    }
    cum += counts.back();
    o += name + "_bucket{" + labels + ",le=\"+Inf\"} " + std::to_string(cum) + "\n";
    snprintf(b, sizeof b, "%.9g", sum);
    o += name + "_sum{" + labels + "} " + b + "\n";
    o += name + "_count{" + labels + "} " + std::to_string(n) + "\n";
  }
};

struct Registry {
  std::string model;
  std::atomic<int64_t> running{0}, waiting{0};
  std::atomic<int64_t> prompt_tokens{0}, generation_tokens{0};
  std::atomic<int64_t> success_stop{0}, success_length{0}, success_abort{0};
  std::atomic<int64_t> kv_used_blocks{0}, kv_total_blocks{1};
  std::atomic<int64_t> prefix_queries{0}, prefix_hits{0};  // prompt tokens looked up in / served from the prefix cache
  Histogram ttft{{0.001, 0.005, 0.01, 0.02, 0.04, 0.06, 0.08, 0.1, 0.25, 0.5, 0.75, 1.0, 2.5, 5.0, 7.5, 10.0, 20.0, 40.0, 80.0, 160.0, 640.0, 2560.0}};
  Histogram itl{{0.01, 0.025, 0.05, 0.075, 0.1, 0.15, 0.2, 0.3, 0.4, 0.5, 0.75, 1.0, 2.5, 5.0, 7.5, 10.0, 20.0, 40.0, 80.0}};
  Histogram e2e{{0.3, 0.5, 0.8, 1.0, 1.5, 2.0, 2.5, 5.0, 10.0, 15.0, 20.0, 30.0, 40.0, 50.0, 60.0, 120.0, 240.0, 480.0, 960.0, 1920.0, 7680.0}};

  std::string render() {
    const std::string L = "model_name=\"" + model + "\"";
    std::string o;
    auto gauge = [&](const std::string& n, const std::string& help, double v) {
      char b[64];
      snprintf(b, sizeof b, "%.9g", v);
      o += "# HELP " + n + " " + help + "\n# TYPE " + n + " gauge\n" + n + "{" + L + "} " + b + "\n";
    };
    auto counter = [&](const std::string& n, const std::string& help, int64_t v, const std::string& extra = "") {
      o += "# HELP " + n + " " + help + "\n# TYPE " + n + " counter\n" + n + "{" + L + extra + "} " + std::to_string(v) + "\n";
    };
    gauge("vllm:num_requests_running", "Number of requests in model execution batches.", (double)running.load());
    gauge("vllm:num_requests_waiting", "Number of requests waiting to be processed.", (double)waiting.load());
    const double usage = (double)kv_used_blocks.load() / (double)std::max<int64_t>(1, kv_total_blocks.load());
    gauge("vllm:kv_cache_usage_perc", "KV-cache usage. 1 means 100 percent usage.", usage);
    gauge("vllm:gpu_cache_usage_perc", "GPU KV-cache usage (legacy name). 1 means 100 percent usage.", usage);
    counter("vllm:prompt_tokens_total", "Number of prefill tokens processed.", prompt_tokens.load());
    counter("vllm:generation_tokens_total", "Number of generation tokens processed.", generation_tokens.load());
    counter("vllm:prefix_cache_queries_total", "Prefix cache queries, in number of queried tokens.", prefix_queries.load());
    counter("vllm:prefix_cache_hits_total", "Prefix cache hits, in number of cached tokens.", prefix_hits.load());
    o += "# HELP vllm:request_success_total Count of successfully processed requests.\n# TYPE vllm:request_success_total counter\n";
    o += "vllm:request_success_total{" + L + ",finished_reason=\"stop\"} " + std::to_string(success_stop.load()) + "\n";
    o += "vllm:request_success_total{" + L + ",finished_reason=\"length\"} " + std::to_string(success_length.load()) + "\n";
    o += "vllm:request_success_total{" + L + ",finished_reason=\"abort\"} " + std::to_string(success_abort.load()) + "\n";
    ttft.render(o, "vllm:time_to_first_token_seconds", "Histogram of time to first token in seconds.", L);
    itl.render(o, "vllm:inter_token_latency_seconds", "Histogram of inter-token latency in seconds.", L);
    itl.render(o, "vllm:time_per_output_token_seconds", "Histogram of time per output token in seconds (legacy name).", L);
    e2e.render(o, "vllm:e2e_request_latency_seconds", "Histogram of end to end request latency in seconds.", L);
    return o;
  }
};

}  // namespace b2metrics
