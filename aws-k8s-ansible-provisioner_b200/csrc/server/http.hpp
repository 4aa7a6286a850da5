// http.hpp -- small blocking HTTP/1.1 server: one thread per connection (the gateway in front of the
// pod keeps a handful of long-lived connections, llm-d-test.yaml:32-43,61-73), keep-alive, fixed-length
// and chunked (SSE) responses.  No TLS: the pod sits behind the cluster's inference gateway.
#pragma once
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cstring>
#include <functional>
#include <map>
#include <string>
#include <thread>

namespace b2http {

struct Request {
  std::string method, path, query, body;
  std::map<std::string, std::string> headers;  // lower-cased names
};

class Conn {
 public:
  explicit Conn(int fd) : fd_(fd) {}
  bool ok() const { return ok_; }
  void respond(int status, const std::string& content_type, const std::string& body, bool keep_alive = true) {
    std::string h = "HTTP/1.1 " + std::to_string(status) + " " + reason(status) + "\r\nContent-Type: " + content_type +
                    "\r\nContent-Length: " + std::to_string(body.size()) + "\r\nConnection: " + (keep_alive ? "keep-alive" : "close") +
                    "\r\n\r\n";
    write_all(h + body);
  }
  // server-sent events: chunked transfer, one chunk per event
  void begin_sse() {
    write_all("HTTP/1.1 200 OK\r\nContent-Type: text/event-stream; charset=utf-8\r\nCache-Control: no-cache\r\n"
              "Connection: keep-alive\r\nTransfer-Encoding: chunked\r\n\r\n");
  }
  void sse_event(const std::string& data) {
    std::string payload = "data: " + data + "\n\n";
    char len[32];
    snprintf(len, sizeof len, "%zx\r\n", payload.size());
    write_all(std::string(len) + payload + "\r\n");
  }
  void end_sse() { write_all("0\r\n\r\n"); }
  // bytes that already are an HTTP response (a proxy relaying a backend's answer)
  void raw_write(const char* p, size_t n) { write_all(std::string(p, n)); }

 private:
  int fd_;
  bool ok_ = true;
  static const char* reason(int s) {
    switch (s) {
      case 200: return "OK"; case 400: return "Bad Request"; case 404: return "Not Found"; case 405: return "Method Not Allowed";
      case 413: return "Payload Too Large"; case 500: return "Internal Server Error"; case 503: return "Service Unavailable";
      default: return "OK";
    }
  }
  void write_all(const std::string& s) {
    size_t off = 0;
    while (ok_ && off < s.size()) {
      ssize_t n = ::send(fd_, s.data() + off, s.size() - off, MSG_NOSIGNAL);
      if (n <= 0) { ok_ = false; break; }
      off += (size_t)n;
    }
  }
};

using Handler = std::function<void(const Request&, Conn&)>;

class Server {
 public:
  Server(std::string host, int port, Handler h) : host_(std::move(host)), port_(port), handler_(std::move(h)) {}
  // returns the bound port (port 0 = pick a free one), or -1
  int listen_and_start() {
    signal(SIGPIPE, SIG_IGN);
    lfd_ = ::socket(AF_INET, SOCK_STREAM, 0);
    if (lfd_ < 0) return -1;
    int one = 1;
    setsockopt(lfd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in a{};
    a.sin_family = AF_INET;
    a.sin_port = htons((uint16_t)port_);
    if (host_.empty() || host_ == "0.0.0.0") a.sin_addr.s_addr = INADDR_ANY;
    else if (inet_pton(AF_INET, host_.c_str(), &a.sin_addr) != 1) return -1;
    if (::bind(lfd_, (sockaddr*)&a, sizeof a) != 0) return -1;
    if (::listen(lfd_, 512) != 0) return -1;
    socklen_t len = sizeof a;
    getsockname(lfd_, (sockaddr*)&a, &len);
    port_ = ntohs(a.sin_port);
    acceptor_ = std::thread([this] { accept_loop(); });
    return port_;
  }
  void stop() {
    stopping_ = true;
    if (lfd_ >= 0) { ::shutdown(lfd_, SHUT_RDWR); ::close(lfd_); lfd_ = -1; }
    if (acceptor_.joinable()) acceptor_.join();
  }
  int port() const { return port_; }

 private:
  std::string host_;
  int port_;
  Handler handler_;
  int lfd_ = -1;
  std::thread acceptor_;
  std::atomic<bool> stopping_{false};
  std::atomic<int> live_{0};

  void accept_loop() {
    while (!stopping_) {
      int fd = ::accept(lfd_, nullptr, nullptr);
      if (fd < 0) { if (stopping_) break; continue; }
      if (live_ >= 2048) { ::close(fd); continue; }
      int one = 1;
      setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
      ++live_;
      std::thread([this, fd] { serve(fd); ::close(fd); --live_; }).detach();
    }
  }
  static std::string lower(std::string s) { for (auto& c : s) c = (char)tolower((unsigned char)c); return s; }
  void serve(int fd) {
    std::string buf;
    char tmp[16384];
    while (!stopping_) {
      size_t hdr_end;
      while ((hdr_end = buf.find("\r\n\r\n")) == std::string::npos) {
        ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
        if (n <= 0) return;
        buf.append(tmp, (size_t)n);
        if (buf.size() > (1u << 20)) return;
      }
      Request req;
      {
        size_t le = buf.find("\r\n");
        std::string line = buf.substr(0, le);
        size_t s1 = line.find(' '), s2 = line.rfind(' ');
        if (s1 == std::string::npos || s2 == s1) return;
        req.method = line.substr(0, s1);
        std::string target = line.substr(s1 + 1, s2 - s1 - 1);
        size_t qm = target.find('?');
        req.path = target.substr(0, qm);
        if (qm != std::string::npos) req.query = target.substr(qm + 1);
        size_t p = le + 2;
        while (p < hdr_end) {
          size_t e = buf.find("\r\n", p);
          std::string h = buf.substr(p, e - p);
          size_t c = h.find(':');
          if (c != std::string::npos) {
            std::string v = h.substr(c + 1);
            size_t a = v.find_first_not_of(" \t");
            req.headers[lower(h.substr(0, c))] = a == std::string::npos ? "" : v.substr(a);
          }
          p = e + 2;
        }
      }
      size_t clen = 0;
      auto it = req.headers.find("content-length");
      if (it != req.headers.end()) clen = (size_t)strtoull(it->second.c_str(), nullptr, 10);
      Conn conn(fd);
      if (clen > (64u << 20)) { conn.respond(413, "application/json", "{\"error\":{\"message\":\"body too large\",\"type\":\"invalid_request_error\",\"code\":413}}", false); return; }
      size_t have = buf.size() - (hdr_end + 4);
      while (have < clen) {
        ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
        if (n <= 0) return;
        buf.append(tmp, (size_t)n);
        have += (size_t)n;
      }
      req.body = buf.substr(hdr_end + 4, clen);
      buf.erase(0, hdr_end + 4 + clen);
      handler_(req, conn);
      if (!conn.ok()) return;
      auto ch = req.headers.find("connection");
      if (ch != req.headers.end() && lower(ch->second) == "close") return;
    }
  }
};

}  // namespace b2http
