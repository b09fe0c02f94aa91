// .svati scene loader behind the reference's parser API
// (gpu/headers/parser.h:18, gpu/headers/parse_obj.h:8-9; behaviour of
// gpu/parser.cpp:6-112 and gpu/parse_obj.cpp:8-62, format in SURVEY Appendix B).
//
// The reference reads tokens with fscanf("%s") / "%f" one at a time.  This loader reads
// the file once and walks it with a pointer tokenizer + strtof (what glibc's %f does),
// which is what SURVEY 8(f)-2 asks for: the 10 M-triangle scene is ~2 GB of text.
// Semantics kept: whitespace-separated tokens, last `camera` wins, lights and objects in
// file order, `#` only at top level, LIFO reversal of v / vn inside an object
// (memory triangle k corner j = file vertex NV-1-(3k+j)), errx(1, ...) on bad input.
#include <cerrno>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <err.h>
#include <string>
#include <vector>
#include "rt_host.h"

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
  const char *end;

  void skip_ws() {
    while (p < end && is_space(*p)) ++p;
  }
  // next whitespace-delimited token (fscanf "%s"); false at EOF
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
    char *stop = nullptr;
    float v = strtof(p, &stop);  // buffer is NUL-terminated by load_file
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

void push_light(struct scene *s, const struct light &l) {
  s->lights = static_cast<struct light *>(realloc(s->lights, sizeof(struct light) * (s->light_count + 1)));
  s->lights[s->light_count++] = l;
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

void read_object(Cursor &c, struct scene *s) {
  long nv_signed = c.integer("object vertex count");
  if (nv_signed < 0) fail("negative vertex count");
  const uint32_t nv = (uint32_t)nv_signed;
  // add_object_to_scene may move objects[]; take the pointer after the call only
  struct object *o = add_object_to_scene(s, nv / 3);

  std::vector<vector3> v, vn;
  v.reserve(nv);
  vn.reserve(nv);
  const char *tok;
  size_t len;
  while ((v.size() < nv || vn.size() < nv) && c.token(tok, len)) {
    if (is_tok(tok, len, "v")) v.push_back(c.vec("v"));
    else if (is_tok(tok, len, "vn")) vn.push_back(c.vec("vn"));
    else if (is_tok(tok, len, "Ka")) o->ka = c.vec("Ka");
    else if (is_tok(tok, len, "Kd")) o->kd = c.vec("Kd");
    else if (is_tok(tok, len, "Ks")) o->ks = c.vec("Ks");
    else if (is_tok(tok, len, "Ns")) o->ns = c.real("Ns");
    else if (is_tok(tok, len, "Ni")) o->ni = c.real("Ni");
    else if (is_tok(tok, len, "Nr")) o->nr = c.real("Nr");
    else if (is_tok(tok, len, "d")) o->d = c.real("d");
    else fail("Error during parsing %.*s\n", (int)len, tok);
  }
  const size_t need = (size_t)o->triangle_count * 3;
  if (v.size() < need || vn.size() < need) fail("object truncated: %zu v / %zu vn for %u vertices", v.size(), vn.size(), nv);
  // LIFO pop order of the reference's two stacks
  for (size_t k = 0; k < need; ++k) {
    o->triangles.vertex[k] = v[v.size() - 1 - k];
    o->triangles.normal[k] = vn[vn.size() - 1 - k];
  }
}

void parse_buffer(const char *data, size_t size, struct scene *s) {
  Cursor c{data, data + size};
  const char *tok;
  size_t len;
  while (c.token(tok, len)) {
    if (is_tok(tok, len, "camera")) {
      struct camera cam;
      cam.width = (int)c.integer("camera width");
      cam.height = (int)c.integer("camera height");
      cam.position = c.vec("camera position");
      cam.u = c.vec("camera u");
      cam.v = c.vec("camera v");
      cam.fov = c.real("camera fov");
      s->camera = cam;
    } else if (is_tok(tok, len, "a_light")) {
      push_light(s, read_light(c, AMBIENT));
    } else if (is_tok(tok, len, "d_light")) {
      push_light(s, read_light(c, DIRECTIONAL));
    } else if (is_tok(tok, len, "p_light")) {
      push_light(s, read_light(c, POINT));
    } else if (is_tok(tok, len, "object")) {
      read_object(c, s);
    } else if (is_tok(tok, len, "#")) {
      c.skip_comment();
    } else {
      fail("Error during the parsing %.*s\n", (int)len, tok);
    }
  }
}

bool load_file(const char *path, std::vector<char> &buf, std::string &why) {
  FILE *f = fopen(path, "rb");
  if (!f) {
    why = strerror(errno);
    return false;
  }
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  buf.resize((size_t)(n > 0 ? n : 0) + 1);
  size_t got = n > 0 ? fread(buf.data(), 1, (size_t)n, f) : 0;
  fclose(f);
  buf[got] = '\0';
  buf.resize(got + 1);
  return true;
}

}  // namespace

extern "C" int rtb_parse_svati(const char *path, struct scene *out, char *errbuf, size_t errlen) {
  *out = empty_scene();
  std::vector<char> buf;
  std::string why;
  if (!load_file(path, buf, why)) {
    if (errbuf && errlen) snprintf(errbuf, errlen, "%s\n", why.c_str());
    return -1;
  }
  try {
    parse_buffer(buf.data(), buf.size() - 1, out);
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
    read_object(c, scene);
  } catch (const ParseError &e) {
    errx(1, "%s", e.msg.c_str());
  }
  if (start >= 0) fseek(file, start + (long)(c.p - buf.data()), SEEK_SET);
}
