// rt_math.hpp -- host-side mirror of the reference's `math` module (src/math/mod.rs:7-11):
// Matrix, Vector3, Point3, Ray, plus scene::Color.  Same names, same argument meaning, same
// f32 arithmetic (this code is compiled with -ffp-contract=off: Rust never fuses a*b+c),
// because the matrices built here ARE the geometry the device traces: inverse() must give
// the bits the reference's Gauss-Jordan gives (src/math/matrix.rs:88-142).
//
// Host only.  Nothing here intersects or shades anything: that exists on the device alone.
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>
#include <stdexcept>

namespace rust_tracer {

constexpr float F32_EPSILON = std::numeric_limits<float>::epsilon();

struct Matrix;

struct Vector3 { // src/math/vector3.rs
    float x = 0.f, y = 0.f, z = 0.f;
    Vector3() = default;
    Vector3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    Vector3 neg() const { return {-x, -y, -z}; }
    Vector3 scalar_mul(float a) const { return {x * a, y * a, z * a}; }
    Vector3 scalar_div(float d) const { return {x / d, y / d, z / d}; }
    Vector3 add(const Vector3 &v) const { return {x + v.x, y + v.y, z + v.z}; }
    Vector3 sub(const Vector3 &v) const { return {x - v.x, y - v.y, z - v.z}; }
    float len2() const { return x * x + y * y + z * z; }
    float len() const { return std::sqrt(len2()); }
    float dot(const Vector3 &v) const { return x * v.x + y * v.y + z * v.z; }
    Vector3 norm() const { return scalar_div(len()); }
    Vector3 cross(const Vector3 &v) const { return {y * v.z - z * v.y, z * v.x - x * v.z, x * v.y - y * v.x}; }
    Vector3 reflect(const Vector3 &about) const { return about.scalar_mul(2.f * dot(about)).sub(*this); }
    bool operator==(const Vector3 &r) const { // vector3.rs:118-125
        return std::fabs(r.x - x) < F32_EPSILON && std::fabs(r.y - y) < F32_EPSILON && std::fabs(r.z - z) < F32_EPSILON;
    }
};
inline Vector3 operator-(const Vector3 &a) { return a.neg(); }
inline Vector3 operator+(const Vector3 &a, const Vector3 &b) { return a.add(b); }
inline Vector3 operator-(const Vector3 &a, const Vector3 &b) { return a.sub(b); }
inline Vector3 operator*(const Vector3 &a, float s) { return a.scalar_mul(s); }
inline Vector3 operator*(float s, const Vector3 &a) { return a.scalar_mul(s); }

struct Point3 { // src/math/point.rs
    float x = 0.f, y = 0.f, z = 0.f;
    Point3() = default;
    Point3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    Point3 scalar_mul(float a) const { return {x * a, y * a, z * a}; }
    Vector3 sub(const Point3 &q) const { return {x - q.x, y - q.y, z - q.z}; }
    Point3 add(const Vector3 &v) const { return {x + v.x, y + v.y, z + v.z}; }
    bool operator==(const Point3 &r) const { return x == r.x && y == r.y && z == r.z; }
};
inline Vector3 operator-(const Point3 &a, const Point3 &b) { return a.sub(b); }
inline Point3 operator+(const Point3 &a, const Vector3 &v) { return a.add(v); }

struct Ray { // src/math/ray.rs
    Point3 origin_;
    Vector3 direction_;
    Ray() = default;
    Ray(const Point3 &o, const Vector3 &d) : origin_(o), direction_(d) {}
    Point3 origin() const { return origin_; }
    Vector3 direction() const { return direction_; }
    Ray neg() const { return {origin_, direction_.neg()}; }
    Ray norm() const { return {origin_, direction_.norm()}; }
    Point3 scalar_mul(float a) const { return origin_ + direction_ * a; }
};
inline Point3 operator*(float t, const Ray &r) { return r.scalar_mul(t); }

struct Matrix { // src/math/matrix.rs -- row major, column-vector convention
    float mat[4][4];

    Matrix() { // Matrix::new(): zeros
        for (auto &row : mat)
            for (float &v : row) v = 0.f;
    }
    float get(int row, int col) const { return mat[row][col]; }
    static Matrix identity() {
        Matrix m;
        for (int i = 0; i < 4; ++i) m.mat[i][i] = 1.f;
        return m;
    }
    static Matrix scale(float x, float y, float z) {
        Matrix m = identity();
        m.mat[0][0] = x; m.mat[1][1] = y; m.mat[2][2] = z;
        return m;
    }
    static Matrix translate(float x, float y, float z) {
        Matrix m = identity();
        m.mat[0][3] = x; m.mat[1][3] = y; m.mat[2][3] = z;
        return m;
    }
    // angles in degrees (matrix.rs:166-206)
    static Matrix rotate_x(float angle) {
        const float rads = to_rads(angle), c = std::cos(rads), s = std::sin(rads);
        Matrix m = identity();
        m.mat[1][1] = c; m.mat[1][2] = -s; m.mat[2][1] = s; m.mat[2][2] = c;
        return m;
    }
    static Matrix rotate_y(float angle) {
        const float rads = to_rads(angle), c = std::cos(rads), s = std::sin(rads);
        Matrix m = identity();
        m.mat[0][0] = c; m.mat[0][2] = s; m.mat[2][0] = -s; m.mat[2][2] = c;
        return m;
    }
    static Matrix rotate_z(float angle) {
        const float rads = to_rads(angle), c = std::cos(rads), s = std::sin(rads);
        Matrix m = identity();
        m.mat[0][0] = c; m.mat[0][1] = -s; m.mat[1][0] = s; m.mat[1][1] = c;
        return m;
    }
    Matrix scalar_mul(float a) const {
        Matrix m;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) m.mat[r][c] = a * mat[r][c];
        return m;
    }
    Matrix scalar_div(float a) const {
        Matrix m;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) m.mat[r][c] = mat[r][c] / a;
        return m;
    }
    Matrix mat_mul(const Matrix &a) const { // matrix.rs:57-75: the sum starts from 0.
        Matrix m;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) {
                float sum = 0.f;
                for (int i = 0; i < 4; ++i) sum += mat[r][i] * a.mat[i][c];
                m.mat[r][c] = sum;
            }
        return m;
    }
    Matrix transpose() const {
        Matrix m;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) m.mat[r][c] = mat[c][r];
        return m;
    }
    // matrix.rs:94-142.  Throws std::runtime_error("Singular Matrix") where the reference panics.
    void invert() {
        Matrix inv = identity();
        for (int col = 0; col < 4; ++col) {
            if (std::fabs(mat[col][col]) < F32_EPSILON) {
                int big = col;
                for (int row = 0; row < 4; ++row)
                    if (std::fabs(mat[row][col]) > std::fabs(mat[big][col])) big = row;
                if (big == col) throw std::runtime_error("Singular Matrix");
                for (int j = 0; j < 4; ++j) {
                    std::swap(mat[big][j], mat[col][j]);
                    std::swap(inv.mat[big][j], inv.mat[col][j]);
                }
            }
            for (int row = 0; row < 4; ++row) {
                if (row == col) continue;
                const float coeff = mat[row][col] / mat[col][col];
                if (std::fabs(coeff) >= F32_EPSILON) {
                    for (int j = 0; j < 4; ++j) {
                        mat[row][j] -= coeff * mat[col][j];
                        inv.mat[row][j] -= coeff * inv.mat[col][j];
                    }
                    mat[row][col] = 0.f;
                }
            }
        }
        for (int row = 0; row < 4; ++row)
            for (int col = 0; col < 4; ++col) inv.mat[row][col] /= mat[row][row];
        *this = inv;
    }
    Matrix inverse() const {
        Matrix copy = *this;
        copy.invert();
        return copy;
    }
    Vector3 vec3_mul(const Vector3 &v) const {
        return {v.x * mat[0][0] + v.y * mat[0][1] + v.z * mat[0][2], v.x * mat[1][0] + v.y * mat[1][1] + v.z * mat[1][2],
                v.x * mat[2][0] + v.y * mat[2][1] + v.z * mat[2][2]};
    }
    Point3 pt_mul(const Point3 &p) const {
        return {p.x * mat[0][0] + p.y * mat[0][1] + p.z * mat[0][2] + mat[0][3],
                p.x * mat[1][0] + p.y * mat[1][1] + p.z * mat[1][2] + mat[1][3],
                p.x * mat[2][0] + p.y * mat[2][1] + p.z * mat[2][2] + mat[2][3]};
    }
    Ray ray_mul(const Ray &r) const { return {pt_mul(r.origin()), vec3_mul(r.direction())}; }
    bool operator==(const Matrix &o) const {
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c)
                if (mat[r][c] != o.mat[r][c]) return false;
        return true;
    }
    void store(float *out16) const {
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) out16[r * 4 + c] = mat[r][c];
    }

  private:
    static float to_rads(float angle) { return angle / 180.0f * 3.14159265358979323846264338327950288f; }
};
inline Matrix operator*(const Matrix &a, const Matrix &b) { return a.mat_mul(b); }
inline Vector3 operator*(const Matrix &m, const Vector3 &v) { return m.vec3_mul(v); }
inline Point3 operator*(const Matrix &m, const Point3 &p) { return m.pt_mul(p); }
inline Ray operator*(const Matrix &m, const Ray &r) { return m.ray_mul(r); }

struct Color { // src/scene/color.rs
    float r = 0.f, g = 0.f, b = 0.f;
    Color() = default;
    Color(float r_, float g_, float b_) : r(r_), g(g_), b(b_) {}
    bool operator==(const Color &o) const {
        return std::fabs(r - o.r) < F32_EPSILON && std::fabs(g - o.g) < F32_EPSILON && std::fabs(b - o.b) < F32_EPSILON;
    }
};
inline Color operator+(const Color &a, const Color &b) { return {a.r + b.r, a.g + b.g, a.b + b.b}; }
inline Color operator*(const Color &a, const Color &b) { return {a.r * b.r, a.g * b.g, a.b * b.b}; }
inline Color operator*(float s, const Color &c) { return {s * c.r, s * c.g, s * c.b}; }
namespace colors {
const Color BLACK(0.f, 0.f, 0.f), WHITE(1.f, 1.f, 1.f), RED(1.f, 0.f, 0.f), GREEN(0.f, 1.f, 0.f), BLUE(0.f, 0.f, 1.f);
}

} // namespace rust_tracer
