// srt_listread.h -- Fortran list-directed numeric input as gfortran's run-time library performs it, host side.
//
// Every text file the reference reads on this path goes through `read(unit,*) list`: the ray file (raytracer_driver.f95:1146),
// the Ngo cards (ngo_dens_model.f95:61-115) and the grid file (interp_dens_model_adapter.f95:59-118).  The rules below are the
// library's, case by case (tests/test_formats_gfortran.py runs the same READ statements on libgfortran itself and compares):
//   * values are separated by blanks (tab, CR and the end of a record count as blanks), by ONE comma or semicolon, or by a slash;
//   * nothing between two separators -- or before the first separator of a READ -- is a null value: the item keeps what it held;
//   * r*c is r copies of c, r* is r null values, 0* is an error;
//   * a slash ends the READ: the remaining items keep what they held;
//   * a READ takes records until its list is served, then drops what is left of the record it ended in;
//   * a value is [sign] digits[.digits] | .digits, then E|D|Q [sign] digits or just sign digits (1.5+3); or [sign] INF |
//     INFINITY | NAN | NAN(...), any case; converted by strtod (correctly rounded; overflow -> Infinity, underflow -> 0 / subnormal);
//   * anything else is an input error (iostat > 0); the end of the file before the list is served is the end condition (iostat < 0).
#pragma once
#include <ctype.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <string>

namespace srt_host {

inline bool fortran_number(const char* tok, size_t n, double& v) {
  size_t i = 0;
  char small[64];
  std::string big;
  char* t = small;
  if (n + 4 > sizeof small) { big.resize(n + 4); t = &big[0]; }  // (a digit string of any length is a legal value)
  size_t m = 0;
  bool neg = false;
  if (i < n && (tok[i] == '+' || tok[i] == '-')) { neg = tok[i] == '-'; t[m++] = tok[i++]; }
  if (i < n && isalpha((unsigned char)tok[i])) {
    std::string w;
    for (size_t k = i; k < n; ++k) w.push_back((char)tolower((unsigned char)tok[k]));
    if (w == "inf" || w == "infinity") { v = neg ? -HUGE_VAL : HUGE_VAL; return true; }
    if (w == "nan" || (w.size() >= 5 && w.compare(0, 4, "nan(") == 0 && w.back() == ')' && w.find(')') == w.size() - 1)) { v = NAN; return true; }
    return false;
  }
  size_t digits = 0;
  while (i < n && isdigit((unsigned char)tok[i])) { t[m++] = tok[i++]; ++digits; }
  if (i < n && tok[i] == '.') {
    t[m++] = tok[i++];
    while (i < n && isdigit((unsigned char)tok[i])) { t[m++] = tok[i++]; ++digits; }
  }
  if (digits == 0) return false;
  if (i < n) {
    const char c = tok[i];
    if (c == 'e' || c == 'E' || c == 'd' || c == 'D' || c == 'q' || c == 'Q') ++i;
    else if (c != '+' && c != '-') return false;
    t[m++] = 'e';
    if (i < n && (tok[i] == '+' || tok[i] == '-')) t[m++] = tok[i++];
    size_t ed = 0;
    while (i < n && isdigit((unsigned char)tok[i])) { t[m++] = tok[i++]; ++ed; }
    if (ed == 0 || i != n) return false;
  }
  t[m] = 0;
  v = strtod(t, nullptr);
  return true;
}
inline bool fortran_number(const std::string& tok, double& v) { return fortran_number(tok.data(), tok.size(), v); }

struct ListReader {
  enum { END_OF_FILE = -1, BAD_VALUE = -2 };
  std::string text;
  size_t p = 0;

  bool open(const char* path) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    char chunk[1 << 16];
    for (size_t got; (got = fread(chunk, 1, sizeof chunk, f)) > 0;) text.append(chunk, got);
    fclose(f);
    p = 0;
    return true;
  }
  static bool is_blank(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }
  static bool is_end(char c) { return is_blank(c) || c == ',' || c == ';' || c == '/'; }

  // One READ statement with `want` numeric items.  v[0..want) is in/out: an item given a null value, or left over after a slash,
  // keeps what v held.  Returns want when the READ completed, END_OF_FILE or BAD_VALUE otherwise (the file position is then
  // undefined, as after a failed READ).
  long read(long want, double* v) {
    long idx = 0;
    bool expecting = true, slash = false;
    const size_t n = text.size();
    while (idx < want && !slash) {
      while (p < n && is_blank(text[p])) ++p;
      if (p >= n) return END_OF_FILE;
      const char c = text[p];
      if (c == ',' || c == ';') { ++p; if (expecting) ++idx; else expecting = true; continue; }
      if (c == '/') { ++p; slash = true; break; }
      size_t q = p;
      while (q < n && !is_end(text[q])) ++q;
      const char* tok = text.data() + p;
      size_t len = q - p;
      p = q;
      long rep = 1;
      size_t d = 0;
      while (d < len && isdigit((unsigned char)tok[d])) ++d;
      if (d > 0 && d < len && tok[d] == '*') {
        rep = d > 15 ? (1L << 60) : atol(std::string(tok, d).c_str());
        if (rep == 0) return BAD_VALUE;
        tok += d + 1;
        len -= d + 1;
        if (len == 0) { idx += rep < want ? rep : want; expecting = false; continue; }  // r*: r null values
      }
      double val;
      if (!fortran_number(tok, len, val)) return BAD_VALUE;
      for (; rep > 0 && idx < want; --rep) v[idx++] = val;
      expecting = false;
    }
    while (p < n && text[p] != '\n') ++p;  // the rest of the record the READ ended in
    if (p < n) ++p;
    return want;
  }
};

}  // namespace srt_host
