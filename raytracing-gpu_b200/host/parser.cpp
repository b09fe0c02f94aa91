// .svati scene loader behind the reference's parser API
// (gpu/headers/parser.h:18, gpu/headers/parse_obj.h:8-9; behaviour of
// gpu/parser.cpp:6-112 and gpu/parse_obj.cpp:8-62, format in SURVEY Appendix B).
//
// The reference reads tokens with fscanf("%s") / "%f" one at a time.  This loader reads the
// file once, cuts it at the top-level `object` tokens and parses the pieces on all host
// threads with a pointer tokenizer and std::from_chars (correctly rounded like glibc's %f;
// strtof for anything from_chars does not take), then assembles the scene in file order:
// SURVEY 8(f)-2, the 10 M-triangle scene is ~2.3 GB of text.
// Semantics kept: whitespace-separated tokens, last `camera` wins, lights and objects in
// file order, `#` only at top level, LIFO reversal of v / vn inside an object
// (memory triangle k corner j = file vertex NV-1-(3k+j)), errx(1, ...) on bad input.  Where a
// cut cannot be proven harmless (a comment running into it) the file is parsed in one piece.
#include <algorithm>
#include <atomic>
#include <cerrno>
#include <charconv>
#include <chrono>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <err.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
#include <string>
#include <thread>
#include <vector>
#include "rt_host.h"

extern "C" int rtb_scene_reserve(struct scene *s, size_t triangles);   // scene.cpp
extern "C" void rtb_scene_mem_pageable(int on);                        // scene.cpp
// Present when this file is part of librtb200.so (csrc/api.cu), absent in the host-only library.
extern "C" void rtb_warmup(void) __attribute__((weak));

namespace {

struct ParseError {
  std::string msg;
};

[[noreturn]] void fail(const char *fmt, ...) {
  char buf[256];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  throw ParseError{buf};
}

bool is_space(char c) { return c == ' ' || c == '\n' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }

struct Cursor {
  const char *p;
  const char *end;          // end of this piece; the whole buffer is NUL-terminated beyond it
  bool comment_cut = false; // a comment ran into `end` without a newline

  void skip_ws() {
    while (p < end && is_space(*p)) ++p;
  }
  // next whitespace-delimited token (fscanf "%s"); false at the end of the piece
  bool token(const char *&tok, size_t &len) {
    skip_ws();
    if (p >= end) return false;
    tok = p;
    while (p < end && !is_space(*p)) ++p;
    len = (size_t)(p - tok);
    return true;
  }
  float real(const char *what) {
    skip_ws();
    if (p >= end) fail("unexpected end of file while reading %s", what);
    // fast path: [+-]digits[.digits][e[+-]digits], what every .svati writer emits
    const char *q = p;
    if (*q == '+' && q + 1 < end && (q[1] == '.' || (q[1] >= '0' && q[1] <= '9'))) ++q;
    float v = 0;
    const std::from_chars_result r = std::from_chars(q, end, v, std::chars_format::general);
    if (r.ec == std::errc() && r.ptr > q) {
      // from_chars and strtof agree on decimal input; hex floats ("0x1p3") are strtof-only
      if (!(r.ptr < end && (*r.ptr == 'x' || *r.ptr == 'X'))) {
        p = r.ptr;
        return v;
      }
    }
    char *stop = nullptr;
    v = strtof(p, &stop);     // inf/nan spellings, hex floats, out-of-range values: glibc's own answer
    if (stop == p) fail("expected a number for %s", what);
    p = stop;
    return v;
  }
  long integer(const char *what) {
    skip_ws();
    if (p >= end) fail("unexpected end of file while reading %s", what);
    char *stop = nullptr;
    long v = strtol(p, &stop, 10);
    if (stop == p) fail("expected an integer for %s", what);
    p = stop;
    return v;
  }
  // fscanf(" %[^\n]"): skip whitespace (newlines included), then the rest of that line
  void skip_comment() {
    skip_ws();
    while (p < end && *p != '\n') ++p;
    if (p >= end) comment_cut = true;
  }
  vector3 vec(const char *what) {
    vector3 v;
    v.x = real(what);
    v.y = real(what);
    v.z = real(what);
    return v;
  }
};

bool is_tok(const char *tok, size_t len, const char *lit) {
  return strlen(lit) == len && memcmp(tok, lit, len) == 0;
}

struct light read_light(Cursor &c, enum light_type type) {
  struct light l;
  memset(&l, 0, sizeof l);
  l.type = type;
  l.r = c.real("light colour");
  l.g = c.real("light colour");
  l.b = c.real("light colour");
  if (type != AMBIENT) l.v = c.vec("light vector");
  return l;
}

// One `object` block, parsed into private storage (file order).
struct ObjectData {
  uint32_t nv = 0;
  std::vector<vector3> v, vn;
  struct object mat;          // material fields only; defaults of add_object_to_scene
  bool has[7] = {false, false, false, false, false, false, false};   // Ka Kd Ks Ns Ni Nr d given in the file
};

// `more_follows`: another piece starts right after this one (with the token `object`)
void read_object(Cursor &c, ObjectData &o, bool more_follows) {
  long nv_signed = c.integer("object vertex count");
  if (nv_signed < 0) fail("negative vertex count");
  o.nv = (uint32_t)nv_signed;
  memset(&o.mat, 0, sizeof o.mat);
  o.v.reserve(o.nv);
  o.vn.reserve(o.nv);
  const char *tok;
  size_t len;
  while (o.v.size() < o.nv || o.vn.size() < o.nv) {
    if (!c.token(tok, len)) {
      // the one-piece parser would read the next token here, which is the `object` of the next piece
      if (more_follows) fail("Error during parsing object\n");
      break;
    }
    if (is_tok(tok, len, "v")) o.v.push_back(c.vec("v"));
    else if (is_tok(tok, len, "vn")) o.vn.push_back(c.vec("vn"));
    else if (is_tok(tok, len, "Ka")) { o.mat.ka = c.vec("Ka"); o.has[0] = true; }
    else if (is_tok(tok, len, "Kd")) { o.mat.kd = c.vec("Kd"); o.has[1] = true; }
    else if (is_tok(tok, len, "Ks")) { o.mat.ks = c.vec("Ks"); o.has[2] = true; }
    else if (is_tok(tok, len, "Ns")) { o.mat.ns = c.real("Ns"); o.has[3] = true; }
    else if (is_tok(tok, len, "Ni")) { o.mat.ni = c.real("Ni"); o.has[4] = true; }
    else if (is_tok(tok, len, "Nr")) { o.mat.nr = c.real("Nr"); o.has[5] = true; }
    else if (is_tok(tok, len, "d")) { o.mat.d = c.real("d"); o.has[6] = true; }
    else fail("Error during parsing %.*s\n", (int)len, tok);
  }
  const size_t need = (size_t)(o.nv / 3) * 3;
  if (o.v.size() < need || o.vn.size() < need)
    fail("object truncated: %zu v / %zu vn for %u vertices", o.v.size(), o.vn.size(), o.nv);
}

// What the top level of a piece contains, in file order.
struct TopItem {
  enum Kind { CAMERA, LIGHT, OBJECT } kind;
  struct camera cam;
  struct light light;
  size_t object;            // index into Piece::objects
};

struct Piece {
  const char *begin = nullptr, *end = nullptr;
  bool more_follows = false;
  std::vector<TopItem> items;
  std::vector<ObjectData> objects;
  bool failed = false;
  std::string error;
  bool comment_cut = false;
};

void parse_piece(Piece &pc) {
  Cursor c{pc.begin, pc.end};
  const char *tok;
  size_t len;
  try {
    while (c.token(tok, len)) {
      TopItem it;
      memset(&it.cam, 0, sizeof it.cam);
      memset(&it.light, 0, sizeof it.light);
      it.object = 0;
      if (is_tok(tok, len, "camera")) {
        it.kind = TopItem::CAMERA;
        it.cam.width = (int)c.integer("camera width");
        it.cam.height = (int)c.integer("camera height");
        it.cam.position = c.vec("camera position");
        it.cam.u = c.vec("camera u");
        it.cam.v = c.vec("camera v");
        it.cam.fov = c.real("camera fov");
        pc.items.push_back(it);
      } else if (is_tok(tok, len, "a_light") || is_tok(tok, len, "d_light") || is_tok(tok, len, "p_light")) {
        it.kind = TopItem::LIGHT;
        it.light = read_light(c, tok[0] == 'a' ? AMBIENT : tok[0] == 'd' ? DIRECTIONAL : POINT);
        pc.items.push_back(it);
      } else if (is_tok(tok, len, "object")) {
        it.kind = TopItem::OBJECT;
        it.object = pc.objects.size();
        pc.objects.emplace_back();
        read_object(c, pc.objects.back(), pc.more_follows);
        pc.items.push_back(it);
      } else if (is_tok(tok, len, "#")) {
        c.skip_comment();
      } else {
        fail("Error during the parsing %.*s\n", (int)len, tok);
      }
    }
  } catch (const ParseError &e) {
    pc.failed = true;
    pc.error = e.msg;
  }
  pc.comment_cut = c.comment_cut && pc.more_follows;
}

// Positions of the tokens `object` (whitespace or buffer start before, whitespace after).
std::vector<const char *> object_tokens(const char *data, size_t size) {
  std::vector<const char *> at;
  const char *p = data, *end = data + size;
  while (p + 6 <= end) {
    const char *q = static_cast<const char *>(memchr(p, 'o', (size_t)(end - p)));
    if (!q || q + 6 > end) break;
    if (memcmp(q, "object", 6) == 0 && (q == data || is_space(q[-1])) && (q + 6 == end || is_space(q[6]))) {
      at.push_back(q);
      p = q + 6;
    } else {
      p = q + 1;
    }
  }
  return at;
}

unsigned host_threads() {
  unsigned n = std::thread::hardware_concurrency();
  if (const char *e = getenv("RTB_PARSE_THREADS")) n = (unsigned)atoi(e);
  return n < 1 ? 1 : (n > 64 ? 64 : n);
}

template <typename F>
void parallel_for(size_t n, unsigned threads, F f) {
  if (threads <= 1 || n <= 1) {
    for (size_t i = 0; i < n; ++i) f(i);
    return;
  }
  std::atomic<size_t> next{0};
  std::vector<std::thread> pool;
  const unsigned t = (unsigned)std::min<size_t>(threads, n);
  for (unsigned k = 0; k < t; ++k)
    pool.emplace_back([&] {
      for (size_t i; (i = next.fetch_add(1)) < n;) f(i);
    });
  for (std::thread &th : pool) th.join();
}

void push_light(struct scene *s, const struct light &l) {
  s->lights = static_cast<struct light *>(realloc(s->lights, sizeof(struct light) * (s->light_count + 1)));
  s->lights[s->light_count++] = l;
}

// ObjectData -> scene: add_object_to_scene, material, LIFO pop order of the reference's two stacks
void place_object(struct scene *s, const ObjectData &od, unsigned threads) {
  // add_object_to_scene may move objects[]; take the pointer after the call only
  struct object *o = add_object_to_scene(s, od.nv / 3);
  if (od.has[0]) o->ka = od.mat.ka;
  if (od.has[1]) o->kd = od.mat.kd;
  if (od.has[2]) o->ks = od.mat.ks;
  if (od.has[3]) o->ns = od.mat.ns;
  if (od.has[4]) o->ni = od.mat.ni;
  if (od.has[5]) o->nr = od.mat.nr;
  if (od.has[6]) o->d = od.mat.d;
  const size_t need = (size_t)o->triangle_count * 3;
  const size_t nvv = od.v.size(), nvn = od.vn.size();
  vector3 *dv = o->triangles.vertex, *dn = o->triangles.normal;
  const size_t chunk = 1 << 18;
  parallel_for((need + chunk - 1) / chunk, need >= 4 * chunk ? threads : 1, [&](size_t ci) {
    const size_t k1 = std::min(need, (ci + 1) * chunk);
    for (size_t k = ci * chunk; k < k1; ++k) {
      dv[k] = od.v[nvv - 1 - k];
      dn[k] = od.vn[nvn - 1 - k];
    }
  });
}

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

void parse_buffer(const char *data, size_t size, struct scene *s) {
  const unsigned threads = host_threads();
  const bool timing = getenv("RTB_PARSE_TIMING") != nullptr;
  const double t0 = now_s();
  std::vector<const char *> cuts = object_tokens(data, size);
  if (size < (1u << 20) || threads <= 1) cuts.clear();          // small files: one piece
  for (int attempt = 0; attempt < 2; ++attempt) {
    // piece 0 = everything before the first `object`; piece i starts at the i-th `object`
    std::vector<Piece> pieces(cuts.size() + 1);
    for (size_t i = 0; i < pieces.size(); ++i) {
      pieces[i].begin = i == 0 ? data : cuts[i - 1];
      pieces[i].end = i < cuts.size() ? cuts[i] : data + size;
      pieces[i].more_follows = i < cuts.size();
    }
    parallel_for(pieces.size(), threads, [&](size_t i) { parse_piece(pieces[i]); });
    const double t1 = now_s();
    // A comment that ran into a cut may have swallowed the `object` after it, and an error
    // next to a cut may read differently: the one-piece parse is the authority for both.
    bool redo = false;
    for (const Piece &pc : pieces) redo |= pc.comment_cut || pc.failed;
    if (redo && !cuts.empty()) {
      cuts.clear();
      continue;
    }
    if (pieces[0].failed) throw ParseError{pieces[0].error};
    size_t total = 0;
    for (const Piece &pc : pieces)
      for (const ObjectData &od : pc.objects) total += od.nv / 3;
    // plain memory: page-locking 720 MB costs more than the one upload of a one-shot run gains
    rtb_scene_mem_pageable(1);
    rtb_scene_reserve(s, total);
    rtb_scene_mem_pageable(0);
    const double t2 = now_s();
    for (Piece &pc : pieces)
      for (const TopItem &it : pc.items) {
        if (it.kind == TopItem::CAMERA) s->camera = it.cam;
        else if (it.kind == TopItem::LIGHT) push_light(s, it.light);
        else {
          place_object(s, pc.objects[it.object], threads);
          std::vector<vector3>().swap(pc.objects[it.object].v);
          std::vector<vector3>().swap(pc.objects[it.object].vn);
        }
      }
    if (timing)
      fprintf(stderr, "parse: %zu pieces on %u threads: tokens %.3f s, reserve %.3f s, assemble %.3f s\n", pieces.size(),
              threads, t1 - t0, t2 - t1, now_s() - t2);
    return;
  }
}

// Whole file into an uninitialised buffer (+ NUL), read with pread on all threads.
struct FileBuf {
  char *data = nullptr;
  size_t size = 0;
  ~FileBuf() { free(data); }
};

bool load_file(const char *path, FileBuf &buf, std::string &why) {
  const int fd = open(path, O_RDONLY);
  if (fd < 0) {
    why = strerror(errno);
    return false;
  }
  struct stat st;
  if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) {
    // not a regular file (pipe, device): plain sequential read
    std::vector<char> acc;
    char chunk[1 << 16];
    ssize_t got;
    while ((got = read(fd, chunk, sizeof chunk)) > 0) acc.insert(acc.end(), chunk, chunk + got);
    close(fd);
    buf.data = static_cast<char *>(malloc(acc.size() + 1));
    if (!buf.data) { why = "out of memory"; return false; }
    memcpy(buf.data, acc.data(), acc.size());
    buf.size = acc.size();
    buf.data[buf.size] = '\0';
    return true;
  }
  const size_t n = (size_t)st.st_size;
  buf.data = static_cast<char *>(malloc(n + 1));
  if (!buf.data) {
    close(fd);
    why = "out of memory";
    return false;
  }
  const size_t chunk = (size_t)8 << 20;
  std::atomic<bool> ok{true};
  parallel_for((n + chunk - 1) / chunk, host_threads(), [&](size_t ci) {
    size_t off = ci * chunk;
    const size_t stop = std::min(n, off + chunk);
    while (off < stop) {
      const ssize_t got = pread(fd, buf.data + off, stop - off, (off_t)off);
      if (got <= 0) { ok = false; return; }
      off += (size_t)got;
    }
  });
  close(fd);
  if (!ok) {
    why = "read error";
    return false;
  }
  buf.size = n;
  buf.data[n] = '\0';
  return true;
}

}  // namespace

extern "C" int rtb_parse_svati(const char *path, struct scene *out, char *errbuf, size_t errlen) {
  *out = empty_scene();
  // the device starts up on a side thread while the file is read and tokenised
  struct Warm {
    std::thread t;
    Warm() { if (rtb_warmup) t = std::thread(rtb_warmup); }
    ~Warm() { if (t.joinable()) t.join(); }
  } warm;
  FileBuf buf;
  std::string why;
  const double t_load = now_s();
  const bool loaded = load_file(path, buf, why);
  if (getenv("RTB_PARSE_TIMING")) fprintf(stderr, "parse: read %zu bytes in %.3f s\n", buf.size, now_s() - t_load);
  if (!loaded) {
    if (errbuf && errlen) snprintf(errbuf, errlen, "%s\n", why.c_str());
    return -1;
  }
  try {
    parse_buffer(buf.data, buf.size, out);
  } catch (const ParseError &e) {
    if (errbuf && errlen) snprintf(errbuf, errlen, "%s", e.msg.c_str());
    free_scene(out);
    return -1;
  }
  return 0;
}

struct scene parser(const char *path) {
  struct scene s;
  char why[256];
  if (rtb_parse_svati(path, &s, why, sizeof why) != 0) errx(1, "%s", why);
  return s;
}

// Stream variants kept for source compatibility with callers of parse_obj.h.
vector3 create_vec(FILE *file) {
  vector3 v = {0, 0, 0};
  if (fscanf(file, "%f %f %f", &v.x, &v.y, &v.z) != 3) errx(1, "Error during parsing: expected 3 floats\n");
  return v;
}

void parse_object(FILE *file, struct scene *scene) {
  // Slurp the remainder of the stream, parse one object block, then seek back to just
  // after it so the caller can go on reading tokens.
  long start = ftell(file);
  std::vector<char> buf;
  char chunk[1 << 16];
  size_t got;
  while ((got = fread(chunk, 1, sizeof chunk, file)) > 0) buf.insert(buf.end(), chunk, chunk + got);
  buf.push_back('\0');
  Cursor c{buf.data(), buf.data() + buf.size() - 1};
  try {
    ObjectData od;
    read_object(c, od, false);
    place_object(scene, od, host_threads());
  } catch (const ParseError &e) {
    errx(1, "%s", e.msg.c_str());
  }
  if (start >= 0) fseek(file, start + (long)(c.p - buf.data()), SEEK_SET);
}
