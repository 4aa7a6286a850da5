// b2route -- load- and prefix-aware HTTP front for N model-server replicas.
//
// In the reference's cluster this job belongs to the llm-d gateway's endpoint picker, which the
// reference's test reaches the model through (llm-d-test.yaml:14-26); on a bare 8-GPU box (BASELINE
// config 5: one b2serve per GPU, ports 8000+i) something has to do it.  The replicas never talk to each
// other (DESIGN.md section 1, row e), so the router is the only place where "which GPU" is decided:
//   * load: every replica's /metrics is scraped (vllm:num_requests_running / _waiting, the same series
//     the endpoint picker reads) and combined with the router's own count of requests in flight;
//   * prefix affinity: requests whose prompt starts with the same text (or ids) go to the replica that
//     served that prefix last, unless it is clearly busier than the least loaded one -- that is what makes
//     a replica's KV/prefix reuse possible at all;
//   * health: a replica whose scrape or connect fails is skipped until a scrape succeeds again.
// The response is relayed byte for byte (Content-Length bodies and chunked/SSE streams alike).
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../json.hpp"
#include "../server/http.hpp"

namespace {

struct Backend {
  std::string host;
  int port = 0;
  std::string name;  // host:port
  std::atomic<int> inflight{0};
  std::atomic<int> running{0}, waiting{0};
  std::atomic<bool> up{false};
  std::atomic<uint64_t> routed{0}, failed{0};
};

struct Router {
  std::vector<std::unique_ptr<Backend>> backends;
  std::string policy = "prefix";  // prefix | least-loaded | round-robin
  int slack = 4;                  // how much busier than the least loaded replica an affinity target may be
  int scrape_ms = 500;
  std::atomic<uint64_t> rr{0}, prefix_hits{0}, prefix_misses{0}, no_backend{0};
  std::mutex mu;
  std::unordered_map<uint64_t, int> affinity;  // prefix hash -> backend index
  std::atomic<bool> stop{false};
};

int connect_to(const std::string& host, int port, int timeout_ms) {
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  if (getaddrinfo(host.c_str(), std::to_string(port).c_str(), &hints, &res) != 0 || !res) return -1;
  int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd >= 0) {
    timeval tv{timeout_ms / 1000, (timeout_ms % 1000) * 1000};
    setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof tv);
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    if (::connect(fd, res->ai_addr, res->ai_addrlen) != 0) {
      ::close(fd);
      fd = -1;
    }
  }
  freeaddrinfo(res);
  return fd;
}

bool send_all(int fd, const std::string& s) {
  size_t off = 0;
  while (off < s.size()) {
    ssize_t n = ::send(fd, s.data() + off, s.size() - off, MSG_NOSIGNAL);
    if (n <= 0) return false;
    off += (size_t)n;
  }
  return true;
}

// Sends `request` to the backend and hands every byte of the response to `sink` as it arrives, stopping at the
// end of the HTTP message (Content-Length or the last chunk), not at connection close.  Returns false if
// nothing could be relayed (the caller may then try another replica).
template <typename Sink>
bool proxy(const Backend& b, const std::string& request, int recv_timeout_s, Sink&& sink, bool* relayed_any) {
  *relayed_any = false;
  int fd = connect_to(b.host, b.port, 2000);
  if (fd < 0) return false;
  timeval tv{recv_timeout_s, 0};
  setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
  bool ok = send_all(fd, request);
  std::string head;
  char tmp[16384];
  size_t hdr_end = std::string::npos;
  while (ok && hdr_end == std::string::npos) {
    ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
    if (n <= 0) { ok = false; break; }
    head.append(tmp, (size_t)n);
    hdr_end = head.find("\r\n\r\n");
    if (head.size() > (1u << 20)) ok = false;
  }
  if (ok) {
    std::string lower = head.substr(0, hdr_end);
    for (auto& c : lower) c = (char)tolower((unsigned char)c);
    const bool chunked = lower.find("transfer-encoding: chunked") != std::string::npos;
    size_t clen = 0;
    const size_t cl = lower.find("content-length:");
    if (cl != std::string::npos) clen = (size_t)strtoull(lower.c_str() + cl + 15, nullptr, 10);
    sink(head.data(), head.size());
    *relayed_any = true;
    if (chunked) {
      // follow the chunk framing only far enough to know where the message ends
      std::string win = head.substr(hdr_end + 4);
      size_t pos = 0;
      bool done = false;
      while (!done) {
        size_t le;
        while (!done && (le = win.find("\r\n", pos)) != std::string::npos) {
          const size_t sz = (size_t)strtoull(win.c_str() + pos, nullptr, 16);
          if (sz == 0) { done = true; break; }
          const size_t next = le + 2 + sz + 2;
          if (next > win.size()) break;
          pos = next;
        }
        if (done) break;
        win.erase(0, pos);
        pos = 0;
        ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
        if (n <= 0) { ok = false; break; }
        sink(tmp, (size_t)n);
        win.append(tmp, (size_t)n);
      }
    } else {
      size_t have = head.size() - (hdr_end + 4);
      while (have < clen) {
        ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
        if (n <= 0) { ok = false; break; }
        sink(tmp, (size_t)n);
        have += (size_t)n;
      }
    }
  }
  ::close(fd);
  return ok;
}

std::string build_request(const b2http::Request& r, const Backend& b) {
  std::string target = r.path + (r.query.empty() ? "" : "?" + r.query);
  std::string s = r.method + " " + target + " HTTP/1.1\r\nHost: " + b.name + "\r\nConnection: close\r\n";
  auto ct = r.headers.find("content-type");
  if (ct != r.headers.end()) s += "Content-Type: " + ct->second + "\r\n";
  auto au = r.headers.find("authorization");
  if (au != r.headers.end()) s += "Authorization: " + au->second + "\r\n";
  auto ac = r.headers.find("accept");
  if (ac != r.headers.end()) s += "Accept: " + ac->second + "\r\n";
  if (r.method == "POST" || !r.body.empty()) s += "Content-Length: " + std::to_string(r.body.size()) + "\r\n";
  return s + "\r\n" + r.body;
}

void scrape_loop(Router* R) {
  while (!R->stop) {
    for (auto& bp : R->backends) {
      Backend& b = *bp;
      std::string body;
      bool any = false;
      const std::string req = "GET /metrics HTTP/1.1\r\nHost: " + b.name + "\r\nConnection: close\r\n\r\n";
      const bool ok = proxy(b, req, 2, [&](const char* p, size_t n) { body.append(p, n); }, &any);
      if (!ok || body.compare(0, 12, "HTTP/1.1 200") != 0) {
        b.up = false;
        continue;
      }
      auto value_of = [&](const char* name) -> int {
        size_t p = 0;
        double total = -1;
        while ((p = body.find(name, p)) != std::string::npos) {
          const bool line_start = p == 0 || body[p - 1] == '\n';
          const size_t eol = body.find('\n', p);
          const std::string line = body.substr(p, eol == std::string::npos ? std::string::npos : eol - p);
          p += strlen(name);
          if (!line_start) continue;
          const size_t sp = line.rfind(' ');
          if (sp == std::string::npos) continue;
          total = (total < 0 ? 0 : total) + atof(line.c_str() + sp + 1);
        }
        return total < 0 ? 0 : (int)total;
      };
      b.running = value_of("vllm:num_requests_running");
      b.waiting = value_of("vllm:num_requests_waiting");
      b.up = true;
    }
    for (int i = 0; i < R->scrape_ms / 20 && !R->stop; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
}

uint64_t fnv1a(const std::string& s) {
  uint64_t h = 1469598103934665603ull;
  for (unsigned char c : s) { h ^= c; h *= 1099511628211ull; }
  return h;
}

// What identifies the START of a request's prompt: the first 256 bytes of the prompt text, the first 64 prompt
// ids, or (chat) the first 256 bytes of the concatenated roles and contents.
bool prefix_key(const b2http::Request& r, uint64_t* key) {
  if (r.body.empty()) return false;
  try {
    b2json::Value j = b2json::parse(r.body);
    std::string k;
    if (const b2json::Value* p = j.find("prompt")) {
      if (p->is_string()) k = "s:" + p->str.substr(0, 256);
      else if (p->is_array() && p->size() > 0) {
        const b2json::Value& first = p->at(0);
        if (first.is_string()) k = "s:" + first.str.substr(0, 256);
        else {
          const b2json::Value& ids = first.is_array() ? first : *p;
          k = "i:";
          for (size_t i = 0; i < ids.size() && i < 64; ++i) k += std::to_string(ids.at(i).as_int()) + ",";
        }
      }
    } else if (const b2json::Value* m = j.find("messages")) {
      k = "m:";
      for (size_t i = 0; i < m->size() && k.size() < 258; ++i) {
        const b2json::Value* role = m->at(i).find("role");
        const b2json::Value* content = m->at(i).find("content");
        k += (role ? role->as_string() : "") + "\x01" + (content ? content->as_string() : "") + "\x02";
      }
      k = k.substr(0, 258);
    }
    if (k.size() <= 2) return false;
    *key = fnv1a(k);
    return true;
  } catch (const std::exception&) {
    return false;
  }
}

int load_of(const Backend& b) { return b.inflight.load() + b.waiting.load(); }

// Picks a replica; `exclude` = one that has just failed for this request (-1 = none).
int pick(Router* R, bool has_key, uint64_t key, int exclude) {
  const int n = (int)R->backends.size();
  int best = -1;
  if (R->policy == "round-robin") {
    for (int k = 0; k < n; ++k) {
      const int i = (int)((R->rr++) % (uint64_t)n);
      if (i != exclude && R->backends[i]->up) return i;
    }
    return -1;
  }
  for (int i = 0; i < n; ++i) {
    if (i == exclude || !R->backends[i]->up) continue;
    if (best < 0 || load_of(*R->backends[i]) < load_of(*R->backends[best])) best = i;
  }
  if (best < 0 || R->policy != "prefix" || !has_key) return best;
  std::lock_guard<std::mutex> l(R->mu);
  auto it = R->affinity.find(key);
  int chosen = best;
  if (it != R->affinity.end() && it->second != exclude && R->backends[it->second]->up &&
      load_of(*R->backends[it->second]) <= load_of(*R->backends[best]) + R->slack) {
    chosen = it->second;
    R->prefix_hits++;
  } else {
    R->prefix_misses++;
  }
  if (R->affinity.size() > 65536) R->affinity.clear();
  R->affinity[key] = chosen;
  return chosen;
}

std::string metrics_text(Router* R) {
  std::string o;
  o += "# HELP b2route_requests_total Requests relayed to each replica.\n# TYPE b2route_requests_total counter\n";
  for (auto& b : R->backends) o += "b2route_requests_total{backend=\"" + b->name + "\"} " + std::to_string(b->routed.load()) + "\n";
  o += "# HELP b2route_backend_failures_total Requests that could not be relayed to the replica first chosen.\n# TYPE b2route_backend_failures_total counter\n";
  for (auto& b : R->backends) o += "b2route_backend_failures_total{backend=\"" + b->name + "\"} " + std::to_string(b->failed.load()) + "\n";
  o += "# HELP b2route_backend_up 1 if the replica's last /metrics scrape succeeded.\n# TYPE b2route_backend_up gauge\n";
  for (auto& b : R->backends) o += "b2route_backend_up{backend=\"" + b->name + "\"} " + std::to_string(b->up ? 1 : 0) + "\n";
  o += "# HELP b2route_backend_load Requests in flight through the router plus the replica's reported queue.\n# TYPE b2route_backend_load gauge\n";
  for (auto& b : R->backends) o += "b2route_backend_load{backend=\"" + b->name + "\"} " + std::to_string(load_of(*b)) + "\n";
  o += "# TYPE b2route_prefix_hits_total counter\nb2route_prefix_hits_total " + std::to_string(R->prefix_hits.load()) + "\n";
  o += "# TYPE b2route_prefix_misses_total counter\nb2route_prefix_misses_total " + std::to_string(R->prefix_misses.load()) + "\n";
  o += "# TYPE b2route_no_backend_total counter\nb2route_no_backend_total " + std::to_string(R->no_backend.load()) + "\n";
  return o;
}

void handle(Router* R, const b2http::Request& r, b2http::Conn& c) {
  if (r.method == "GET" && r.path == "/metrics") { c.respond(200, "text/plain; version=0.0.4", metrics_text(R)); return; }
  if (r.method == "GET" && r.path == "/health") {
    for (auto& b : R->backends)
      if (b->up) { c.respond(200, "text/plain", ""); return; }
    c.respond(503, "text/plain", "no replica is up");
    return;
  }
  uint64_t key = 0;
  const bool has_key = r.method == "POST" && prefix_key(r, &key);
  int exclude = -1;
  for (int attempt = 0; attempt < 2; ++attempt) {
    const int i = pick(R, has_key, key, exclude);
    if (i < 0) break;
    Backend& b = *R->backends[i];
    b.inflight++;
    bool any = false;
    const bool ok = proxy(b, build_request(r, b), 600, [&](const char* p, size_t n) { c.raw_write(p, n); }, &any);
    b.inflight--;
    if (ok || any) {  // once bytes have gone to the client the request cannot be replayed elsewhere
      b.routed++;
      return;
    }
    b.failed++;
    b.up = false;  // until the next successful scrape
    exclude = i;
  }
  R->no_backend++;
  c.respond(503, "application/json", "{\"error\":{\"message\":\"no model-server replica available\",\"type\":\"service_unavailable\",\"code\":503}}");
}

}  // namespace

int main(int argc, char** argv) {
  Router R;
  std::string host = "0.0.0.0", backends;
  int port = 8080;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto val = [&]() -> std::string { return i + 1 < argc ? argv[++i] : ""; };
    if (a == "--port") port = atoi(val().c_str());
    else if (a == "--host") host = val();
    else if (a == "--backends") backends = val();
    else if (a == "--policy") R.policy = val();
    else if (a == "--slack") R.slack = atoi(val().c_str());
    else if (a == "--scrape-interval-ms") R.scrape_ms = atoi(val().c_str());
    else {
      fprintf(stderr, "usage: b2route --backends host:port[,host:port...] [--port 8080] [--host H] [--policy prefix|least-loaded|round-robin]\n"
                      "               [--slack 4] [--scrape-interval-ms 500]\n");
      return a == "--help" ? 0 : 2;
    }
  }
  size_t p = 0;
  while (p < backends.size()) {
    size_t e = backends.find(',', p);
    if (e == std::string::npos) e = backends.size();
    const std::string item = backends.substr(p, e - p);
    const size_t colon = item.rfind(':');
    if (colon != std::string::npos) {
      auto b = std::make_unique<Backend>();
      b->host = item.substr(0, colon);
      b->port = atoi(item.c_str() + colon + 1);
      b->name = item;
      R.backends.push_back(std::move(b));
    }
    p = e + 1;
  }
  if (R.backends.empty() || (R.policy != "prefix" && R.policy != "least-loaded" && R.policy != "round-robin")) {
    fprintf(stderr, "b2route: need --backends and a known --policy\n");
    return 2;
  }
  std::thread scraper(scrape_loop, &R);
  b2http::Server srv(host, port, [&R](const b2http::Request& r, b2http::Conn& c) { handle(&R, r, c); });
  const int bound = srv.listen_and_start();
  if (bound < 0) {
    fprintf(stderr, "b2route: cannot listen on %s:%d\n", host.c_str(), port);
    R.stop = true;
    scraper.join();
    return 1;
  }
  printf("READY %d\n", bound);
  fflush(stdout);
  while (true) std::this_thread::sleep_for(std::chrono::seconds(3600));
  return 0;
}
