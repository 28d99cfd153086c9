// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the shipped library.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build, load or call anything under oracle/.
//
// Minimal GLSL vector/builtin shim so that the reference's shader bodies can be
// restated in C++ nearly token for token (no glm in this image).
//
// PARITY UNPINNED: the reference has no tests, golden vectors or CPU code for this
// path and its shaders cannot be compiled or run here (no GLSL compiler, no Vulkan).
// Where GLSL leaves arithmetic implementation-defined this shim takes a documented
// position (all fp32, round-to-nearest-even, no FMA contraction; build with
// -ffp-contract=off):
//   min(x,y)      = y < x ? y : x                (GLSL 4.50 §8.3)
//   max(x,y)      = x < y ? y : x
//   clamp(x,a,b)  = min(max(x,a),b)
//   mix(x,y,a)    = x*(1-a) + y*a
//   step(e,x)     = x < e ? 0 : 1
//   sign(x)       = x>0 ? 1 : x<0 ? -1 : 0
//   fract(x)      = x - floor(x)
//   smoothstep    = t*t*(3-2t), t = clamp((x-e0)/(e1-e0),0,1)
//   dot           = left-to-right sum of products
//   length(v)     = sqrt(dot(v,v));  normalize(v) = v / length(v)
//   cross         = GLSL 4.50 §8.5 component formulas
//   pow(x,1.0)    = x   (voronoise is only ever called with v = 1.0 → k = 1.0)
//   float→int     = truncate, saturating, NaN→0   (CUDA cvt.rzi semantics)
//   float→uint    = truncate, saturating at 0 and 2^32-1, NaN→0
//   pow(x, n)     = x*x*...*x for the integer exponents 3, 5, 20 of the shading code
//   pow(b, e)     = (float)exp((double)e * log((double)b)) for the one real exponent (frag.glsl:703)
#pragma once
#include <cmath>
#include <cstdint>

namespace glsl {

typedef uint32_t uint;

inline int f2i(float f) {
    if (f != f) return 0;
    if (f >= 2147483648.0f) return 2147483647;
    if (f <= -2147483648.0f) return (int)(-2147483647 - 1);
    return (int)f;
}
inline uint f2u(float f) {
    if (f != f) return 0u;
    if (f <= 0.0f) return 0u;
    if (f >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint)f;
}

struct vec2 { float x, y;
    vec2() : x(0), y(0) {}
    explicit vec2(float s) : x(s), y(s) {}
    vec2(float a, float b) : x(a), y(b) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};
struct vec3 { float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    vec3(vec2 v, float c) : x(v.x), y(v.y), z(c) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
    vec2 xz() const { return vec2(x, z); }
    vec2 xy() const { return vec2(x, y); }
    vec3 yzx() const { return vec3(y, z, x); }
    vec3 zxy() const { return vec3(z, x, y); }
};
struct vec4 { float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
    vec3 xyz() const { return vec3(x, y, z); }
    void set_xyz(vec3 v) { x = v.x; y = v.y; z = v.z; }
};
struct ivec3 { int x, y, z;
    ivec3() : x(0), y(0), z(0) {}
    explicit ivec3(int s) : x(s), y(s), z(s) {}
    ivec3(int a, int b, int c) : x(a), y(b), z(c) {}
    explicit ivec3(vec3 v) : x(f2i(v.x)), y(f2i(v.y)), z(f2i(v.z)) {}
    int& operator[](int i) { return (&x)[i]; }
    int operator[](int i) const { return (&x)[i]; }
};
struct uvec2 { uint x, y;
    uvec2() : x(0), y(0) {}
    uvec2(uint a, uint b) : x(a), y(b) {}
};
struct uvec3 { uint x, y, z;
    uvec3() : x(0), y(0), z(0) {}
    explicit uvec3(uint s) : x(s), y(s), z(s) {}
    uvec3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
    explicit uvec3(vec3 v) : x(f2u(v.x)), y(f2u(v.y)), z(f2u(v.z)) {}
    uvec3(ivec3 v) : x((uint)v.x), y((uint)v.y), z((uint)v.z) {}  // GLSL implicit int→uint
    uint& operator[](int i) { return (&x)[i]; }
    uint operator[](int i) const { return (&x)[i]; }
};
struct ivec4 { int x, y, z, w;
    ivec4() : x(0), y(0), z(0), w(0) {}
    ivec4(int a, int b, int c, int d) : x(a), y(b), z(c), w(d) {}
    ivec4(vec3 v, int d) : x(f2i(v.x)), y(f2i(v.y)), z(f2i(v.z)), w(d) {}
    ivec3 xyz() const { return ivec3(x, y, z); }
};
struct uvec4 { uint x, y, z, w;
    uvec4() : x(0), y(0), z(0), w(0) {}
    uvec4(uint a, uint b, uint c, uint d) : x(a), y(b), z(c), w(d) {}
    uvec4(uvec3 v, uint d) : x(v.x), y(v.y), z(v.z), w(d) {}
    uvec3 xyz() const { return uvec3(x, y, z); }
};

inline vec3 to_vec3(uvec3 v) { return vec3((float)v.x, (float)v.y, (float)v.z); }
inline vec3 to_vec3(ivec3 v) { return vec3((float)v.x, (float)v.y, (float)v.z); }
inline ivec3 to_ivec3(uvec3 v) { return ivec3((int)v.x, (int)v.y, (int)v.z); }

// ---- float vector arithmetic (component-wise) ----
#define GLSL_VOP(T, N, op) \
    inline T operator op(T a, T b) { T r; for (int i = 0; i < N; i++) r[i] = a[i] op b[i]; return r; } \
    inline T operator op(T a, float b) { T r; for (int i = 0; i < N; i++) r[i] = a[i] op b; return r; } \
    inline T operator op(float a, T b) { T r; for (int i = 0; i < N; i++) r[i] = a op b[i]; return r; }
GLSL_VOP(vec2, 2, +) GLSL_VOP(vec2, 2, -) GLSL_VOP(vec2, 2, *) GLSL_VOP(vec2, 2, /)
GLSL_VOP(vec3, 3, +) GLSL_VOP(vec3, 3, -) GLSL_VOP(vec3, 3, *) GLSL_VOP(vec3, 3, /)
GLSL_VOP(vec4, 4, +) GLSL_VOP(vec4, 4, -) GLSL_VOP(vec4, 4, *) GLSL_VOP(vec4, 4, /)
#undef GLSL_VOP
inline vec2 operator-(vec2 a) { return vec2(-a.x, -a.y); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec4 operator-(vec4 a) { return vec4(-a.x, -a.y, -a.z, -a.w); }
inline vec2& operator+=(vec2& a, vec2 b) { a = a + b; return a; }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
inline vec3& operator-=(vec3& a, vec3 b) { a = a - b; return a; }
inline vec4& operator+=(vec4& a, vec4 b) { a = a + b; return a; }
inline vec4& operator*=(vec4& a, float b) { a = a * b; return a; }
inline bool operator==(vec3 a, vec3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

// ---- integer vector arithmetic ----
inline ivec3 operator+(ivec3 a, ivec3 b) { return ivec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline ivec3 operator-(ivec3 a, ivec3 b) { return ivec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline ivec3 operator*(ivec3 a, ivec3 b) { return ivec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline ivec3 operator*(int a, ivec3 b) { return ivec3(a * b.x, a * b.y, a * b.z); }
inline uvec3 operator+(uvec3 a, uvec3 b) { return uvec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline uvec3 operator-(uvec3 a, uvec3 b) { return uvec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline uvec3 operator*(uvec3 a, uvec3 b) { return uvec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline uvec3 operator*(uint a, uvec3 b) { return uvec3(a * b.x, a * b.y, a * b.z); }
inline uvec3 operator>>(uvec3 a, uint s) { return uvec3(a.x >> s, a.y >> s, a.z >> s); }
inline uvec3 operator&(uvec3 a, uint m) { return uvec3(a.x & m, a.y & m, a.z & m); }
inline uvec3 operator%(uvec3 a, uvec3 b) { return uvec3(a.x % b.x, a.y % b.y, a.z % b.z); }
inline uvec3 operator%(uvec3 a, uint b) { return uvec3(a.x % b, a.y % b, a.z % b); }
inline uvec3 operator/(uvec3 a, uint b) { return uvec3(a.x / b, a.y / b, a.z / b); }

// ---- scalar builtins ----
inline float floor(float x) { return ::floorf(x); }
inline float ceil(float x) { return ::ceilf(x); }
inline float sqrt(float x) { return ::sqrtf(x); }
inline float abs(float x) { return ::fabsf(x); }
inline int iabs(int x) { return x < 0 ? (int)(0u - (uint)x) : x; }  // GLSL abs(int); renamed: ::abs(int) is global in C++
inline float fract(float x) { return x - ::floorf(x); }
inline float min(float x, float y) { return y < x ? y : x; }
inline float max(float x, float y) { return x < y ? y : x; }
inline uint min(uint x, uint y) { return y < x ? y : x; }
inline float clamp(float x, float a, float b) { return min(max(x, a), b); }
inline float mix(float x, float y, float a) { return x * (1.0f - a) + y * a; }
inline float step(float edge, float x) { return x < edge ? 0.0f : 1.0f; }
inline float sign(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
inline float smoothstep(float e0, float e1, float x) {
    float t = clamp((x - e0) / (e1 - e0), 0.0f, 1.0f);
    return t * t * (3.0f - 2.0f * t);
}
inline bool isnan(float x) { return x != x; }

// ---- vector builtins ----
inline vec2 floor(vec2 v) { return vec2(floor(v.x), floor(v.y)); }
inline vec3 floor(vec3 v) { return vec3(floor(v.x), floor(v.y), floor(v.z)); }
inline vec3 ceil(vec3 v) { return vec3(ceil(v.x), ceil(v.y), ceil(v.z)); }
inline vec4 ceil(vec4 v) { return vec4(ceil(v.x), ceil(v.y), ceil(v.z), ceil(v.w)); }
inline vec2 fract(vec2 v) { return vec2(fract(v.x), fract(v.y)); }
inline vec3 fract(vec3 v) { return vec3(fract(v.x), fract(v.y), fract(v.z)); }
inline vec2 abs(vec2 v) { return vec2(abs(v.x), abs(v.y)); }
inline vec3 abs(vec3 v) { return vec3(abs(v.x), abs(v.y), abs(v.z)); }
inline vec3 sign(vec3 v) { return vec3(sign(v.x), sign(v.y), sign(v.z)); }
inline vec2 max(vec2 a, float b) { return vec2(max(a.x, b), max(a.y, b)); }
inline vec3 max(vec3 a, vec3 b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline vec3 min(vec3 a, vec3 b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec2 step(float e, vec2 v) { return vec2(step(e, v.x), step(e, v.y)); }
inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(vec4 a, vec4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline float length(vec2 v) { return sqrt(dot(v, v)); }
inline float length(vec3 v) { return sqrt(dot(v, v)); }
inline float length(vec4 v) { return sqrt(dot(v, v)); }
inline vec2 normalize(vec2 v) { return v / length(v); }
inline vec3 normalize(vec3 v) { return v / length(v); }
inline vec4 normalize(vec4 v) { return v / length(v); }
inline vec3 cross(vec3 x, vec3 y) {
    return vec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
inline vec3 mix(vec3 x, vec3 y, float a) { return x * (1.0f - a) + y * a; }
inline vec3 reflect(vec3 I, vec3 N) { return I - 2.0f * dot(N, I) * N; }  // GLSL 4.50 8.5
inline vec3& operator*=(vec3& a, float b) { a = a * b; return a; }
// pow with a small positive integer exponent: repeated multiplication, left to right (shim position;
// a GLSL pow(x, 20.0) is exp2(20 * log2(x)) on real drivers and undefined for x < 0)
inline float ipow(float x, int n) { float r = x; for (int i = 1; i < n; i++) r = r * x; return r; }
// pow with a real exponent (frag.glsl:703 only): evaluated in double, rounded to float (shim position)
inline float rpow(float b, float e) { return (float)::exp((double)e * ::log((double)b)); }
inline uint bitfieldExtract(uint value, int offset, int bits) {
    return (value >> offset) & ((1u << bits) - 1u);
}
inline uint bitfieldInsert(uint base, uint insert, int offset, int bits) {
    uint mask = ((1u << bits) - 1u) << offset;
    return (base & ~mask) | ((insert << offset) & mask);
}

}  // namespace glsl
