// textio.cu — host-side text formats on either side of the path; no device code.
//   * graph_embs.txt as np.savetxt(fname, hidden_emb) writes it (train.py:193): '%.18e', single space,
//     '\n', no header.  numpy formats every value with Python's '%' operator, i.e. C printf("%.18e") of the
//     value widened to double — reproduced byte for byte, rows formatted in parallel.
//   * node2vec's .embs.txt as train.py:79-80 reads it (header line skipped, first column = node name, the rest
//     parsed by float() = correctly rounded strtod), lines parsed in parallel.
//   * whole_graph.weighted.edgelist as nx.write_weighted_edgelist writes it (predict_drug.py:225-226,
//     run_covid.py:150-151): one line "u v w" per edge with w = str(float) = Python's repr, i.e. the shortest
//     digit string that round-trips, in fixed notation for decimal exponents in [-4, 16), else 'd.ddde+XX'.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <charconv>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"

namespace gdr {

static inline int fmt_e18(char *dst, double v) {
  if (isnan(v)) {  // Python prints 'nan' whatever the sign bit; glibc may print '-nan'
    memcpy(dst, "nan", 3);
    return 3;
  }
  return snprintf(dst, 32, "%.18e", v);
}

template <typename T>
static void format_rows(const T *x, int64_t r0, int64_t r1, int64_t d, std::string *out) {
  out->clear();
  out->reserve((size_t)((r1 - r0) * d * 26));
  char buf[40];
  for (int64_t r = r0; r < r1; ++r) {
    for (int64_t c = 0; c < d; ++c) {
      const int len = fmt_e18(buf, (double)x[r * d + c]);
      out->append(buf, (size_t)len);
      out->push_back(c + 1 == d ? '\n' : ' ');
    }
    if (d == 0) out->push_back('\n');
  }
}

template <typename T>
static int64_t format_all(const T *x, int64_t n, int64_t d, char *out, int64_t cap, int n_threads) {
  if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  if ((int64_t)n_threads > n) n_threads = (int)(n > 0 ? n : 1);
  std::vector<std::string> parts((size_t)n_threads);
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; ++t) {
    const int64_t r0 = n * t / n_threads, r1 = n * (t + 1) / n_threads;
    th.emplace_back(format_rows<T>, x, r0, r1, d, &parts[(size_t)t]);
  }
  for (auto &t : th) t.join();
  int64_t total = 0;
  for (auto &p : parts) total += (int64_t)p.size();
  if (total > cap) return -total;
  int64_t off = 0;
  for (auto &p : parts) {
    memcpy(out + off, p.data(), p.size());
    off += (int64_t)p.size();
  }
  return total;
}

// Python's repr(float) (float_repr_style 'short'): shortest round-trip digits (std::to_chars yields the same digit
// string: both are exact shortest-representation algorithms), laid out by PyOS_double_to_string's 'r' rule.
static int fmt_py_repr(char *dst, double v) {
  if (isnan(v)) { memcpy(dst, "nan", 3); return 3; }
  if (isinf(v)) { const int k = v < 0 ? 4 : 3; memcpy(dst, v < 0 ? "-inf" : "inf", (size_t)k); return k; }
  char sci[40];
  auto r = std::to_chars(sci, sci + sizeof(sci) - 1, v, std::chars_format::scientific);  // [-]d[.ddd]e[+-]XX
  *r.ptr = 0;
  const char *p = sci, *e = r.ptr;
  char *o = dst;
  if (*p == '-') *o++ = *p++;
  char digits[24];
  int nd = 0;
  const char *q = p;
  for (; q < e && *q != 'e'; ++q)
    if (*q != '.') digits[nd++] = *q;
  const int exp10 = atoi(q + 1);           // value = d.ddd x 10^exp10
  const int decpt = exp10 + 1;             // digits to the left of the decimal point
  if (decpt > 16 || decpt < -3) {          // exponent form: d[.ddd]e[+-]XX (at least two exponent digits)
    *o++ = digits[0];
    if (nd > 1) {
      *o++ = '.';
      memcpy(o, digits + 1, (size_t)(nd - 1));
      o += nd - 1;
    }
    o += snprintf(o, 8, "e%c%02d", exp10 < 0 ? '-' : '+', exp10 < 0 ? -exp10 : exp10);
  } else if (decpt <= 0) {                 // 0.000ddd
    *o++ = '0';
    *o++ = '.';
    for (int i = 0; i < -decpt; ++i) *o++ = '0';
    memcpy(o, digits, (size_t)nd);
    o += nd;
  } else if (decpt >= nd) {                // ddd000.0
    memcpy(o, digits, (size_t)nd);
    o += nd;
    for (int i = nd; i < decpt; ++i) *o++ = '0';
    *o++ = '.';
    *o++ = '0';
  } else {                                 // dd.ddd
    memcpy(o, digits, (size_t)decpt);
    o += decpt;
    *o++ = '.';
    memcpy(o, digits + decpt, (size_t)(nd - decpt));
    o += nd - decpt;
  }
  return (int)(o - dst);
}

static int clamp_threads(int n_threads, int64_t work) {
  if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  if ((int64_t)n_threads > work) n_threads = (int)(work > 0 ? work : 1);
  return n_threads;
}

}  // namespace gdr

using namespace gdr;

extern "C" {

// "<name[src]> <name[dst]> <repr(w)>\n" per edge, as nx.write_weighted_edgelist (predict_drug.py:225-226).
// names: one blob + per-node (offset, length).  Returns bytes written, -needed if out_cap is too small (call with
// out == NULL / out_cap == 0 to size the buffer), or a negative GDR_E_* code.
int64_t gdr_text_format_edgelist(const char *names, const int64_t *name_off, const int32_t *name_len, int64_t n_nodes,
                                 const int32_t *src, const int32_t *dst, const double *w, int64_t n_edges, char *out,
                                 int64_t out_cap, int n_threads) {
  GDR_CHECK_ARG(n_nodes >= 0 && n_edges >= 0 && (n_edges == 0 || (names && name_off && name_len && src && dst && w)),
                GDR_E_BADARG, "text_format_edgelist: bad args");
  const int nt = clamp_threads(n_threads, n_edges);
  std::vector<std::string> parts((size_t)nt);
  std::vector<int> bad((size_t)nt, 0);
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t) {
    th.emplace_back([&, t]() {
      std::string &o = parts[(size_t)t];
      char buf[48];
      for (int64_t i = n_edges * t / nt; i < n_edges * (t + 1) / nt; ++i) {
        if (src[i] < 0 || src[i] >= n_nodes || dst[i] < 0 || dst[i] >= n_nodes) { bad[(size_t)t] = 1; return; }
        o.append(names + name_off[src[i]], (size_t)name_len[src[i]]);
        o.push_back(' ');
        o.append(names + name_off[dst[i]], (size_t)name_len[dst[i]]);
        o.push_back(' ');
        o.append(buf, (size_t)fmt_py_repr(buf, w[i]));
        o.push_back('\n');
      }
    });
  }
  for (auto &t : th) t.join();
  for (int b : bad) GDR_CHECK_ARG(!b, GDR_E_BADARG, "text_format_edgelist: node index out of range");
  int64_t total = 0;
  for (auto &p : parts) total += (int64_t)p.size();
  if (!out || total > out_cap) return -total;
  int64_t off = 0;
  for (auto &p : parts) {
    memcpy(out + off, p.data(), p.size());
    off += (int64_t)p.size();
  }
  return total;
}

// nx.read_weighted_edgelist semantics (run_covid.py:150-151 reads what predict_drug.py:225-226 wrote): lines starting
// with '#' and blank lines are skipped, "u v w" split on whitespace, w = float().  Pass 1 (w_out == NULL): returns
// the number of edges.  Pass 2: byte spans of both names and the weight of every edge.  Negative code on a bad line.
int64_t gdr_text_parse_edgelist(const char *text, int64_t len, int64_t *u_off, int32_t *u_len, int64_t *v_off,
                                int32_t *v_len, double *w_out, int n_threads) {
  GDR_CHECK_ARG(text && len >= 0, GDR_E_BADARG, "text_parse_edgelist: bad args");
  auto is_sp = [](char c) { return c == ' ' || c == '\t' || c == '\r'; };
  std::vector<int64_t> starts;
  int64_t pos = 0;
  while (pos < len) {
    const char *nl = static_cast<const char *>(memchr(text + pos, '\n', (size_t)(len - pos)));
    const int64_t end = nl ? (int64_t)(nl - text) : len;
    int64_t i = pos;
    while (i < end && is_sp(text[i])) ++i;
    if (i < end && text[i] != '#') starts.push_back(i);
    pos = end + 1;
  }
  const int64_t n = (int64_t)starts.size();
  if (!w_out) return n;
  GDR_CHECK_ARG(u_off && u_len && v_off && v_len, GDR_E_BADARG, "text_parse_edgelist: missing outputs");
  const int nt = clamp_threads(n_threads, n);
  std::vector<int> bad((size_t)nt, 0);
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t) {
    th.emplace_back([&, t]() {
      std::string tok;
      for (int64_t i = n * t / nt; i < n * (t + 1) / nt; ++i) {
        int64_t p = starts[(size_t)i];
        const char *nl = static_cast<const char *>(memchr(text + p, '\n', (size_t)(len - p)));
        const int64_t e = nl ? (int64_t)(nl - text) : len;
        int64_t span[3][2];
        int k = 0;
        while (p < e && k < 3) {
          while (p < e && is_sp(text[p])) ++p;
          if (p >= e) break;
          span[k][0] = p;
          while (p < e && !is_sp(text[p])) ++p;
          span[k][1] = p;
          ++k;
        }
        while (p < e && is_sp(text[p])) ++p;
        if (k != 3 || p < e) { bad[(size_t)t] = 1; return; }      // networkx raises on a line without exactly 3 fields
        u_off[i] = span[0][0]; u_len[i] = (int32_t)(span[0][1] - span[0][0]);
        v_off[i] = span[1][0]; v_len[i] = (int32_t)(span[1][1] - span[1][0]);
        tok.assign(text + span[2][0], (size_t)(span[2][1] - span[2][0]));
        char *endp = nullptr;
        w_out[i] = strtod(tok.c_str(), &endp);
        if (endp != tok.c_str() + tok.size()) { bad[(size_t)t] = 2; return; }
      }
    });
  }
  for (auto &t : th) t.join();
  for (int b : bad) {
    GDR_CHECK_ARG(b != 1, GDR_E_SHAPE, "text_parse_edgelist: a line does not have exactly 3 fields");
    GDR_CHECK_ARG(b != 2, GDR_E_BADARG, "text_parse_edgelist: a weight is not a number");
  }
  return n;
}

int64_t gdr_text_savetxt_bytes(int64_t n, int64_t d) { return (n < 0 || d < 0) ? GDR_E_BADARG : n * (d * 26 + 1) + 16; }

// bytes written (>= 0), or a negative GDR_E_* code.  x: HOST array [n, d], dtype 4 (fp32) or 8 (fp64) bytes.
int64_t gdr_text_savetxt_e18(const void *x, int64_t n, int64_t d, int elem_bytes, char *out, int64_t out_cap,
                             int n_threads) {
  GDR_CHECK_ARG(n >= 0 && d >= 0 && (n * d == 0 || x) && out, GDR_E_BADARG, "text_savetxt: bad args");
  GDR_CHECK_ARG(elem_bytes == 4 || elem_bytes == 8, GDR_E_SHAPE, "text_savetxt: fp32 or fp64 only");
  const int64_t w = elem_bytes == 4 ? format_all(static_cast<const float *>(x), n, d, out, out_cap, n_threads)
                                    : format_all(static_cast<const double *>(x), n, d, out, out_cap, n_threads);
  GDR_CHECK_ARG(w >= 0, GDR_E_WORKSPACE, "text_savetxt: output buffer too small (%lld needed)", (long long)-w);
  return w;
}

// node2vec .embs.txt: `text` = whole file.  Pass 1 (values == NULL): returns the number of data lines and sets
// *d_out to the number of value columns of the first one.  Pass 2: fills values[N, d] (fp64) and, per line, the
// byte span of the node name (name_off[i], name_len[i]).  Returns N or a negative code (ragged line, bad number).
int64_t gdr_text_parse_embs(const char *text, int64_t len, int skip_rows, double *values, int64_t d,
                            int64_t *name_off, int32_t *name_len, int64_t *d_out, int n_threads) {
  GDR_CHECK_ARG(text && len >= 0 && skip_rows >= 0, GDR_E_BADARG, "text_parse_embs: bad args");
  std::vector<int64_t> starts;
  int64_t pos = 0;
  int skipped = 0;
  while (pos < len) {
    const char *nl = static_cast<const char *>(memchr(text + pos, '\n', (size_t)(len - pos)));
    const int64_t end = nl ? (int64_t)(nl - text) : len;
    bool blank = true;
    for (int64_t i = pos; i < end && blank; ++i) blank = (text[i] == ' ' || text[i] == '\t' || text[i] == '\r');
    if (skipped < skip_rows) ++skipped;
    else if (!blank) starts.push_back(pos);
    pos = end + 1;
  }
  const int64_t n = (int64_t)starts.size();
  auto line_end = [&](int64_t i) {
    const char *nl = static_cast<const char *>(memchr(text + starts[(size_t)i], '\n', (size_t)(len - starts[(size_t)i])));
    return nl ? (int64_t)(nl - text) : len;
  };
  auto is_sp = [](char c) { return c == ' ' || c == '\t' || c == '\r'; };
  if (!values) {
    int64_t cols = 0;
    if (n > 0) {
      const int64_t e = line_end(0);
      int64_t i = starts[0];
      while (i < e) {
        while (i < e && is_sp(text[i])) ++i;
        if (i >= e) break;
        ++cols;
        while (i < e && !is_sp(text[i])) ++i;
      }
    }
    if (d_out) *d_out = cols > 0 ? cols - 1 : 0;
    return n;
  }
  GDR_CHECK_ARG(d >= 0 && name_off && name_len, GDR_E_BADARG, "text_parse_embs: missing outputs");
  if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  std::vector<int> bad((size_t)n_threads, 0);
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; ++t) {
    th.emplace_back([&, t]() {
      std::string tok;
      for (int64_t i = n * t / n_threads; i < n * (t + 1) / n_threads; ++i) {
        const int64_t e = line_end(i);
        int64_t p = starts[(size_t)i];
        while (p < e && is_sp(text[p])) ++p;
        const int64_t n0 = p;
        while (p < e && !is_sp(text[p])) ++p;
        name_off[i] = n0;
        name_len[i] = (int32_t)(p - n0);
        int64_t c = 0;
        while (p < e) {
          while (p < e && is_sp(text[p])) ++p;
          if (p >= e) break;
          const int64_t t0 = p;
          while (p < e && !is_sp(text[p])) ++p;
          if (c >= d) { bad[(size_t)t] = 1; break; }
          tok.assign(text + t0, (size_t)(p - t0));
          char *endp = nullptr;
          const double v = strtod(tok.c_str(), &endp);
          if (endp != tok.c_str() + tok.size()) { bad[(size_t)t] = 2; break; }
          values[i * d + c++] = v;
        }
        if (c != d && !bad[(size_t)t]) bad[(size_t)t] = 1;
      }
    });
  }
  for (auto &t : th) t.join();
  for (int b : bad) {
    GDR_CHECK_ARG(b != 1, GDR_E_SHAPE, "text_parse_embs: a line does not have %lld values", (long long)d);
    GDR_CHECK_ARG(b != 2, GDR_E_BADARG, "text_parse_embs: a token is not a number");
  }
  return n;
}

}  // extern "C"
