// rt_host.cpp — host-side maths of the path (no CUDA): camera basis and per-frame instance
// transforms.  In the reference this code runs on the CPU inside the frame loop, so it stays on
// the CPU here and must reproduce the reference's host arithmetic bit for bit (checked in
// tests/test_host_math.py against values produced by the reference's own compiled header).
// Build with -ffp-contract=off.
#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstring>
#include <string>

#include "../../include/rtb200.h"

namespace {

constexpr float kZero = 1e-6f;       // include/Global/DeviceFunctions.cuh:18
constexpr float kPi = 3.1415926;     // include/Global/DeviceFunctions.cuh:20 (float from a double literal)

struct Vec {
    float x, y, z;
};
Vec sub(Vec a, Vec b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
Vec crossp(Vec a, Vec b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
// normalize() as executed by HOST code in the reference (DeviceFunctions.cuh:397-404): rsqrtf on
// the host is (float)(1.0 / sqrt((double)x)) — CUDA 12.9 crt/math_functions.hpp:2953,3333.
Vec unit_host(Vec a) {
    const float l2 = a.x * a.x + a.y * a.y + a.z * a.z;
    if (l2 <= kZero * kZero) return {0.0f, 0.0f, 1.0f};
    const float inv = static_cast<float>(1.0 / std::sqrt(static_cast<double>(l2)));
    return {a.x * inv, a.y * inv, a.z * inv};
}

// 4x4 product with the reference's accumulation order (Matrix::operator*, DeviceFunctions.cuh:48-61):
// sum starts at 0 and adds a[i][n]*b[n][j] for n ascending.
struct Mat4 {
    float v[4][4];
};
Mat4 matmul(const Mat4 &a, const Mat4 &b) {
    Mat4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float acc = 0.0f;
            for (int n = 0; n < 4; n++) acc += a.v[i][n] * b.v[n][j];
            r.v[i][j] = acc;
        }
    return r;
}
Mat4 identity4() {
    Mat4 m;
    std::memset(&m, 0, sizeof m);
    m.v[0][0] = m.v[1][1] = m.v[2][2] = m.v[3][3] = 1.0f;
    return m;
}
// constructRotateMatrix(degree, axis), DeviceFunctions.cuh:88-121
Mat4 axis_rotation(float degree, int axis) {
    const float theta = degree * kPi / 180.0f;
    const float c = std::cos(theta), s = std::sin(theta);
    Mat4 m = identity4();
    const int a = (axis + 1) % 3, b = (axis + 2) % 3;
    // x: rows/cols (1,2); y: (2,0); z: (0,1) with the sign pattern of a right-handed rotation
    m.v[a][a] = c;
    m.v[a][b] = -s;
    m.v[b][a] = s;
    m.v[b][b] = c;
    return m;
}

// slerp, src/Global/RendererTime.cu:297-340.  Quaternions are float4 {x,y,z,w}; the reference
// returns {w-expression, x-, y-, z-} positionally into {x,y,z,w} — reproduced.
void slerp_ref(const float a[4], const float bin[4], float t, float out[4]) {
    float b[4] = {bin[0], bin[1], bin[2], bin[3]};
    float dp = a[3] * b[3] + a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
    if (dp < 0.0f) {
        b[3] = -b[3]; b[0] = -b[0]; b[1] = -b[1]; b[2] = -b[2];
        dp = -dp;
    }
    if (dp > 0.9995f) {
        float rx = a[3] + t * (b[3] - a[3]);
        float ry = a[0] + t * (b[0] - a[0]);
        float rz = a[1] + t * (b[1] - a[1]);
        float rw = a[2] + t * (b[2] - a[2]);
        const float mag = std::sqrt(rw * rw + rx * rx + ry * ry + rz * rz);
        if (mag > 0.0) {
            rw /= mag; rx /= mag; ry /= mag; rz /= mag;
        }
        out[0] = rx; out[1] = ry; out[2] = rz; out[3] = rw;
        return;
    }
    const float theta0 = std::acos(dp);
    const float theta = theta0 * t;
    const float st = std::sin(theta), st0 = std::sin(theta0);
    const float k0 = std::cos(theta) - dp * st / st0;
    const float k1 = st / st0;
    out[0] = (k0 * a[3]) + (k1 * b[3]);
    out[1] = (k0 * a[0]) + (k1 * b[0]);
    out[2] = (k0 * a[1]) + (k1 * b[1]);
    out[3] = (k0 * a[2]) + (k1 * b[2]);
}

// quatToEuler, src/Global/RendererTime.cu:343-370 (degrees)
void euler_ref(const float q[4], float e[3]) {
    const float qx = q[0], qy = q[1], qz = q[2], qw = q[3];
    const float roll = std::atan2(2 * (qw * qx + qy * qz), 1 - 2 * (qx * qx + qy * qy));
    const float sp = 2 * (qw * qy - qz * qx);
    const float pitch = (std::abs(sp) >= 1) ? std::copysign(kPi / 2, sp) : std::asin(sp);
    const float yaw = std::atan2(2 * (qw * qz + qx * qy), 1 - 2 * (qy * qy + qz * qz));
    e[0] = roll * 180.0f / kPi;
    e[1] = pitch * 180.0f / kPi;
    e[2] = yaw * 180.0f / kPi;
}

// ---- particle colour ramps: include/Util/ColorRamp.cuh:31-112 (stop tables :33-80, bake :83-112)
struct Stop {
    float pos, r, g, b;
};
struct Ramp {
    const char *name;
    int n;
    Stop s[6];
};
const Ramp kRamps[] = {
    {"viridis", 5, {{0.00f, 0.267f, 0.004f, 0.329f}, {0.25f, 0.283f, 0.141f, 0.458f}, {0.50f, 0.254f, 0.265f, 0.530f},
                    {0.75f, 0.196f, 0.509f, 0.364f}, {1.00f, 0.993f, 0.906f, 0.144f}}},
    {"plasma", 5, {{0.00f, 0.050f, 0.030f, 0.527f}, {0.25f, 0.537f, 0.062f, 0.549f}, {0.50f, 0.871f, 0.191f, 0.494f},
                   {0.75f, 0.992f, 0.580f, 0.288f}, {1.00f, 0.940f, 0.975f, 0.131f}}},
    {"spectral", 6, {{0.00f, 0.619f, 0.003f, 0.258f}, {0.20f, 0.835f, 0.243f, 0.310f}, {0.40f, 0.957f, 0.427f, 0.263f},
                     {0.60f, 0.993f, 0.681f, 0.380f}, {0.80f, 0.741f, 0.858f, 0.407f}, {1.00f, 0.400f, 0.761f, 0.647f}}},
    {"terrain", 5, {{0.00f, 0.149f, 0.149f, 0.149f}, {0.25f, 0.114f, 0.451f, 0.208f}, {0.50f, 0.639f, 0.784f, 0.325f},
                    {0.75f, 0.988f, 0.972f, 0.745f}, {1.00f, 0.996f, 0.922f, 0.545f}}},
    {"heatmap", 5, {{0.00f, 0.050f, 0.050f, 0.300f}, {0.25f, 0.000f, 0.000f, 1.000f}, {0.50f, 0.000f, 1.000f, 1.000f},
                    {0.75f, 1.000f, 1.000f, 0.000f}, {1.00f, 1.000f, 0.000f, 0.000f}}},
    {"grayscale", 2, {{0.00f, 0.050f, 0.050f, 0.050f}, {1.00f, 0.950f, 0.950f, 0.950f}}},
};

}  // namespace

extern "C" {

void rt_camera_configure(const float center[3], const float target[3], const float up[3], int flip_up, rt_camera *out) {
    Vec u = unit_host({up[0], up[1], up[2]});
    if (flip_up) u = {-u.x, -u.y, -u.z};
    const Vec c = {center[0], center[1], center[2]};
    const Vec W = sub({target[0], target[1], target[2]}, c);
    const Vec U = unit_host(crossp(W, u));
    const Vec V = unit_host(crossp(U, W));
    const float pack[12] = {c.x, c.y, c.z, U.x, U.y, U.z, V.x, V.y, V.z, W.x, W.y, W.z};
    std::memcpy(out->center, pack, 12);
    std::memcpy(out->U, pack + 3, 12);
    std::memcpy(out->V, pack + 6, 12);
    std::memcpy(out->W, pack + 9, 12);
}

void rt_construct_transform(const float shift[3], const float rot[3], const float scale[3], float out12[12]) {
    Mat4 T = identity4();
    T.v[0][3] = shift[0]; T.v[1][3] = shift[1]; T.v[2][3] = shift[2];
    const Mat4 R = matmul(matmul(axis_rotation(rot[0], 0), axis_rotation(rot[1], 1)), axis_rotation(rot[2], 2));
    Mat4 S = identity4();
    S.v[0][0] = scale[0]; S.v[1][1] = scale[1]; S.v[2][2] = scale[2];
    const Mat4 M = matmul(matmul(T, R), S);
    std::memcpy(out12, &M.v[0][0], 12 * sizeof(float));
}

void rt_transform_time(const float pos[3], const float vel[3], const float qc[4], const float qn[4], float duration,
                       size_t frames, size_t frame, const float pshift[3], const float pscale[3], float out12[12]) {
    float factor = 1.0f;
    if (frames > 1) factor = static_cast<float>(frame) / static_cast<float>(frames - 1);
    float shift[3];
    for (int a = 0; a < 3; a++) {
        const float total = vel[a] * duration;
        const float per_frame = total / static_cast<float>(frames);
        const float here = pos[a] + per_frame * static_cast<float>(frame);
        shift[a] = pshift[a] + here;
    }
    float q[4], e[3];
    slerp_ref(qc, qn, factor, q);
    euler_ref(q, e);
    rt_construct_transform(shift, e, pscale, out12);
}

void rt_transform_mesh(const float vel[3], float duration, size_t frames, size_t frame, const float pshift[3],
                       const float pscale[3], float out12[12]) {
    float shift[3];
    for (int a = 0; a < 3; a++) {
        const float total = vel[a] * duration;
        const float per_frame = total / static_cast<float>(frames);
        shift[a] = pshift[a] + per_frame * static_cast<float>(frame);
    }
    const float zero[3] = {0.0f, 0.0f, 0.0f};
    rt_construct_transform(shift, zero, pscale, out12);
}

// resolveColorRampPreset + colorStopsForPreset + bakeColorRamp, include/Util/ColorRamp.cuh:31-152.
// Unknown names fall back to viridis (:145-151).  out_rgb = float[3*count].  Returns the preset index.
int rt_bake_color_ramp(const char *preset, size_t count, float *out_rgb) {
    std::string key = preset ? preset : "";
    std::transform(key.begin(), key.end(), key.begin(), [](unsigned char c) { return (char)std::tolower(c); });
    int which = 0;
    for (int i = 0; i < (int)(sizeof kRamps / sizeof kRamps[0]); i++)
        if (key == kRamps[i].name) which = i;
    const Ramp &rp = kRamps[which];
    if (count == 0 || !out_rgb) return which;
    if (count == 1) {
        const Stop &b = rp.s[rp.n - 1];
        out_rgb[0] = b.r; out_rgb[1] = b.g; out_rgb[2] = b.b;
        return which;
    }
    for (size_t i = 0; i < count; ++i) {
        const float u = static_cast<float>(i) / static_cast<float>(count - 1);
        const Stop *lower = &rp.s[0];
        const Stop *upper = &rp.s[rp.n - 1];
        for (int k = 1; k < rp.n; ++k) {
            if (u <= rp.s[k].pos) {
                upper = &rp.s[k];
                lower = &rp.s[k - 1];
                break;
            }
            lower = &rp.s[k];
        }
        const float span = upper->pos - lower->pos;
        float t = span > 0.0f ? (u - lower->pos) / span : 0.0f;
        t = std::clamp(t, 0.0f, 1.0f);
        out_rgb[3 * i + 0] = lower->r + (upper->r - lower->r) * t;
        out_rgb[3 * i + 1] = lower->g + (upper->g - lower->g) * t;
        out_rgb[3 * i + 2] = lower->b + (upper->b - lower->b) * t;
    }
    return which;
}

}  // extern "C"
