// TEST INFRASTRUCTURE (oracle shim) -- not product code.
// Minimal stand-in for htslib 1.22 htslib/kstring.h (subprojects/htslib.wrap, un-vendored).
// Call sites: reference gclib/GSam.h:180,543-560; gclib/GSam.cpp:375-385.
#ifndef BRB_SHIM_KSTRING_H
#define BRB_SHIM_KSTRING_H
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <stdio.h>

#ifndef kroundup32
#define kroundup32(x) (--(x), (x)|=(x)>>1, (x)|=(x)>>2, (x)|=(x)>>4, (x)|=(x)>>8, (x)|=(x)>>16, ++(x))
#endif

typedef struct kstring_t { size_t l, m; char *s; } kstring_t;
#define KS_INITIALIZE { 0, 0, NULL }

static inline int ks_resize(kstring_t *s, size_t size) {
  if (s->m < size) {
    size_t m = size < 16 ? 16 : size * 2;
    char *t = (char *)realloc(s->s, m);
    if (!t) return -1;
    s->s = t; s->m = m;
  }
  return 0;
}
static inline int kputsn(const char *p, size_t l, kstring_t *s) {
  if (ks_resize(s, s->l + l + 2) < 0) return -1;
  memcpy(s->s + s->l, p, l); s->l += l; s->s[s->l] = 0; return (int)l;
}
static inline int kputs(const char *p, kstring_t *s) { return kputsn(p, strlen(p), s); }
static inline int kputc(int c, kstring_t *s) {
  if (ks_resize(s, s->l + 2) < 0) return -1;
  s->s[s->l++] = (char)c; s->s[s->l] = 0; return c;
}
static inline int kputw(int c, kstring_t *s) {
  char buf[16]; int n = snprintf(buf, sizeof buf, "%d", c); return kputsn(buf, (size_t)n, s);
}
static inline void ks_free(kstring_t *s) { if (s) { free(s->s); s->s = NULL; s->l = s->m = 0; } }
#endif
