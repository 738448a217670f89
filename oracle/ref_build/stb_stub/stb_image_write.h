// stub: PNG writing is I/O, not part of the path the oracle pins
#pragma once
static inline int stbi_write_png(const char *, int, int, int, const void *, int) { return 0; }
