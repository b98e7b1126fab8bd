// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product: nothing under
// rust_tracer_b200/ may include, link or call this.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs use it, as the checker.
//
// Scalar f32 CPU restatement of the reference's math layer, one IEEE rounding per
// operation in the reference's operand order.  Must be compiled with
// -ffp-contract=off (Rust never fuses a*b+c) and without -ffast-math.
//
// Parity status: the reference is Rust and cannot be built here (no rustc/cargo), and it
// ships no golden images; this restatement is pinned against every known-answer unit test
// the reference holds for this path (tests/test_oracle_kat.py lists them with file:line).
// Whole-image parity is therefore "pinned by KATs + line-by-line review"; a second,
// independent restatement of the same Rust (tests/pyref.py, scalar numpy.float32) agrees with this
// one bit for bit on whole frames, counts and the default scene's description (tests/test_pyref_cpu.py).
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>

namespace orc {

constexpr float EPS = std::numeric_limits<float>::epsilon();  // std::f32::EPSILON
constexpr float PI_F = 3.14159265358979323846264338327950288f; // std::f32::consts::PI

struct Matrix;

// src/math/vector3.rs:6-115
struct Vector3 {
    float x, y, z;
    Vector3() : x(0.f), y(0.f), z(0.f) {}
    Vector3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    Vector3 neg() const { return Vector3(-x, -y, -z); }                        // :30-36
    Vector3 scalar_mul(float a) const { return Vector3(x * a, y * a, z * a); } // :39-45
    Vector3 scalar_div(float d) const { return Vector3(x / d, y / d, z / d); } // :48-54
    Vector3 add(const Vector3 &v) const { return Vector3(x + v.x, y + v.y, z + v.z); }
    Vector3 sub(const Vector3 &v) const { return Vector3(x - v.x, y - v.y, z - v.z); }
    float len2() const { return x * x + y * y + z * z; }                       // :75-77
    float len() const { return std::sqrt(len2()); }                            // :80-83
    float dot(const Vector3 &v) const { return x * v.x + y * v.y + z * v.z; }  // :86-88
    Vector3 norm() const { return scalar_div(len()); }                         // :91-94
    Vector3 cross(const Vector3 &v) const {                                    // :97-103
        return Vector3(y * v.z - z * v.y, z * v.x - x * v.z, x * v.y - y * v.x);
    }
    // :113-115  `2. * (self.dot(about)) * about - self`
    Vector3 reflect(const Vector3 &about) const {
        float s = 2.f * dot(about);
        return about.scalar_mul(s).sub(*this);
    }
};
inline Vector3 operator-(const Vector3 &a) { return a.neg(); }
inline Vector3 operator+(const Vector3 &a, const Vector3 &b) { return a.add(b); }
inline Vector3 operator-(const Vector3 &a, const Vector3 &b) { return a.sub(b); }
inline Vector3 operator*(const Vector3 &a, float s) { return a.scalar_mul(s); }
inline Vector3 operator*(float s, const Vector3 &a) { return a.scalar_mul(s); }

// src/math/point.rs:6-49
struct Point3 {
    float x, y, z;
    Point3() : x(0.f), y(0.f), z(0.f) {}
    Point3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    Vector3 sub(const Point3 &q) const { return Vector3(x - q.x, y - q.y, z - q.z); } // :39-41
    Point3 add(const Vector3 &v) const { return Point3(x + v.x, y + v.y, z + v.z); }  // :43-49
};
inline Vector3 operator-(const Point3 &a, const Point3 &b) { return a.sub(b); }
inline Point3 operator+(const Point3 &a, const Vector3 &v) { return a.add(v); }
inline Vector3 to_vec(const Point3 &p) { return Vector3(p.x, p.y, p.z); } // vector3.rs:128-136

// src/math/ray.rs:7-52
struct Ray {
    Point3 origin;
    Vector3 direction;
    Ray() {}
    Ray(const Point3 &o, const Vector3 &d) : origin(o), direction(d) {}
    // :50-52  `self.origin + self.direction * a`   (f32 * Ray, :63-77)
    Point3 at(float a) const { return origin + direction * a; }
};

// src/math/matrix.rs:5-256 -- row-major, column-vector convention
struct Matrix {
    float mat[4][4];

    static Matrix zero() {
        Matrix m;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) m.mat[r][c] = 0.f;
        return m;
    }
    static Matrix identity() { // :19-28
        Matrix m = zero();
        m.mat[0][0] = m.mat[1][1] = m.mat[2][2] = m.mat[3][3] = 1.f;
        return m;
    }
    static Matrix scale(float x, float y, float z) { // :144-153
        Matrix m = identity();
        m.mat[0][0] = x;
        m.mat[1][1] = y;
        m.mat[2][2] = z;
        return m;
    }
    static Matrix translate(float x, float y, float z) { // :155-164
        Matrix m = identity();
        m.mat[0][3] = x;
        m.mat[1][3] = y;
        m.mat[2][3] = z;
        return m;
    }
    static float rads_of(float angle) { return angle / 180.0f * PI_F; } // :167-168
    static Matrix rotate_x(float angle) {                               // :166-178
        float r = rads_of(angle);
        Matrix m = identity();
        m.mat[1][1] = std::cos(r);
        m.mat[1][2] = -std::sin(r);
        m.mat[2][1] = std::sin(r);
        m.mat[2][2] = std::cos(r);
        return m;
    }
    static Matrix rotate_y(float angle) { // :180-192
        float r = rads_of(angle);
        Matrix m = identity();
        m.mat[0][0] = std::cos(r);
        m.mat[0][2] = std::sin(r);
        m.mat[2][0] = -std::sin(r);
        m.mat[2][2] = std::cos(r);
        return m;
    }
    static Matrix rotate_z(float angle) { // :194-206
        float r = rads_of(angle);
        Matrix m = identity();
        m.mat[0][0] = std::cos(r);
        m.mat[0][1] = -std::sin(r);
        m.mat[1][0] = std::sin(r);
        m.mat[1][1] = std::cos(r);
        return m;
    }

    Matrix mat_mul(const Matrix &a) const { // :57-75  sum starts at 0., i = 0..3
        Matrix out;
        for (int row = 0; row < 4; ++row)
            for (int col = 0; col < 4; ++col) {
                float sum = 0.f;
                for (int i = 0; i < 4; ++i) sum += mat[row][i] * a.mat[i][col];
                out.mat[row][col] = sum;
            }
        return out;
    }
    Matrix transpose() const { // :77-86
        Matrix out;
        for (int row = 0; row < 4; ++row)
            for (int col = 0; col < 4; ++col) out.mat[row][col] = mat[col][row];
        return out;
    }
    // :94-142  Gauss-Jordan; pivots only when |diag| < EPS, skips |coeff| < EPS.
    // Returns false where the reference panics "Singular Matrix".
    bool invert() {
        Matrix inverse = identity();
        for (int col = 0; col < 4; ++col) {
            if (std::fabs(mat[col][col]) < EPS) {
                int big_row = col;
                for (int row = 0; row < 4; ++row)
                    if (std::fabs(mat[row][col]) > std::fabs(mat[big_row][col])) big_row = row;
                if (big_row == col) return false;
                for (int j = 0; j < 4; ++j) {
                    float tmp = mat[big_row][j];
                    mat[big_row][j] = mat[col][j];
                    mat[col][j] = tmp;
                    tmp = inverse.mat[big_row][j];
                    inverse.mat[big_row][j] = inverse.mat[col][j];
                    inverse.mat[col][j] = tmp;
                }
            }
            for (int row = 0; row < 4; ++row) {
                if (row != col) {
                    float coeff = mat[row][col] / mat[col][col];
                    if (std::fabs(coeff) >= EPS) {
                        for (int j = 0; j < 4; ++j) {
                            mat[row][j] -= coeff * mat[col][j];
                            inverse.mat[row][j] -= coeff * inverse.mat[col][j];
                        }
                        mat[row][col] = 0.f;
                    }
                }
            }
        }
        for (int row = 0; row < 4; ++row)
            for (int col = 0; col < 4; ++col) inverse.mat[row][col] /= mat[row][row];
        *this = inverse;
        return true;
    }
    Matrix inverse(bool *ok = nullptr) const { // :88-92
        Matrix copy = *this;
        bool good = copy.invert();
        if (ok) *ok = good;
        return copy;
    }
    Vector3 vec3_mul(const Vector3 &v) const { // :229-235
        return Vector3(v.x * mat[0][0] + v.y * mat[0][1] + v.z * mat[0][2],
                       v.x * mat[1][0] + v.y * mat[1][1] + v.z * mat[1][2],
                       v.x * mat[2][0] + v.y * mat[2][1] + v.z * mat[2][2]);
    }
    Point3 pt_mul(const Point3 &p) const { // :237-252
        return Point3(p.x * mat[0][0] + p.y * mat[0][1] + p.z * mat[0][2] + mat[0][3],
                      p.x * mat[1][0] + p.y * mat[1][1] + p.z * mat[1][2] + mat[1][3],
                      p.x * mat[2][0] + p.y * mat[2][1] + p.z * mat[2][2] + mat[2][3]);
    }
    Ray ray_mul(const Ray &r) const { // :254-256
        return Ray(pt_mul(r.origin), vec3_mul(r.direction));
    }
};
inline Matrix operator*(const Matrix &a, const Matrix &b) { return a.mat_mul(b); } // :259-265

// src/scene/color.rs:1-167
struct Color {
    float r, g, b;
    Color() : r(0.f), g(0.f), b(0.f) {}
    Color(float r_, float g_, float b_) : r(r_), g(g_), b(b_) {}
};
inline Color operator+(const Color &a, const Color &c) { return Color(a.r + c.r, a.g + c.g, a.b + c.b); }
inline Color operator*(const Color &a, const Color &c) { return Color(a.r * c.r, a.g * c.g, a.b * c.b); }
inline Color operator*(float s, const Color &c) { return Color(s * c.r, s * c.g, s * c.b); }
const Color BLACK(0.f, 0.f, 0.f), WHITE(1.f, 1.f, 1.f);

// `(255. * c) as u8` (src/scene/color.rs:43-46): Rust >= 1.45 float->int casts saturate
// and map NaN to 0.
inline uint8_t as_u8_channel(float c) {
    float x = 255.f * c;
    if (!(x == x)) return 0;
    if (x <= 0.f) return 0;
    if (x >= 255.f) return 255;
    return (uint8_t)x; // truncates toward zero
}

} // namespace orc
