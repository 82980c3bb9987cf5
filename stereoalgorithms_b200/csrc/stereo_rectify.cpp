// Rectification transforms from the calibrated pair.
//
// The reference's calibration tool gets R_L, R_R, P1, P2, Q and the valid ROIs from
//   stereoRectify(K_l, D_l, K_r, D_r, size, R, T, R_L, R_R, P1, P2, Q, CALIB_ZERO_DISPARITY, 0, size, &roiL, &roiR)
// (Stereo_Calibration/Stereo_Calibration.cpp:162) and stores them in the YAML file every model wrapper reads
// (FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:98-114).  OpenCV is not part of the reference tree; this file
// restates cv::stereoRectify (Bouguet's algorithm as OpenCV 4.13 implements it) in plain f64 host code so that a
// calibration can be re-rectified for another output size, alpha or flag without OpenCV.  Pinned against
// cv2.stereoRectify 4.13.0 on 48 configurations (tests/golden/stereo_rectify.json, generated with cv2):
// R1, R2 to 1e-15, P1, P2, Q to 1e-12 relative, ROIs equal except where alpha = 0 puts an edge exactly on the image
// border (a ceil() of +-1e-13).
//
// Host-side, per calibration, not per frame: no CUDA here.
#include <cfloat>
#include <cmath>
#include <cstring>

#include "../../include/stereo_geom.h"

namespace {

struct M3 { double m[3][3]; };

M3 mul(const M3& a, const M3& b)
{
    M3 c;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double s = 0.0;
            for (int k = 0; k < 3; k++) s += a.m[i][k] * b.m[k][j];
            c.m[i][j] = s;
        }
    return c;
}

M3 transpose(const M3& a)
{
    M3 c;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) c.m[i][j] = a.m[j][i];
    return c;
}

void mul_vec(const M3& a, const double v[3], double out[3])
{
    for (int i = 0; i < 3; i++) out[i] = a.m[i][0] * v[0] + a.m[i][1] * v[1] + a.m[i][2] * v[2];
}

bool inverse(const M3& a, M3& inv)
{
    const double (*m)[3] = a.m;
    const double c00 = m[1][1] * m[2][2] - m[1][2] * m[2][1], c01 = m[1][2] * m[2][0] - m[1][0] * m[2][2], c02 = m[1][0] * m[2][1] - m[1][1] * m[2][0];
    const double det = m[0][0] * c00 + m[0][1] * c01 + m[0][2] * c02;
    if (det == 0.0 || !std::isfinite(det)) return false;
    const double id = 1.0 / det;
    inv.m[0][0] = c00 * id; inv.m[1][0] = c01 * id; inv.m[2][0] = c02 * id;
    inv.m[0][1] = (m[0][2] * m[2][1] - m[0][1] * m[2][2]) * id;
    inv.m[1][1] = (m[0][0] * m[2][2] - m[0][2] * m[2][0]) * id;
    inv.m[2][1] = (m[0][1] * m[2][0] - m[0][0] * m[2][1]) * id;
    inv.m[0][2] = (m[0][1] * m[1][2] - m[0][2] * m[1][1]) * id;
    inv.m[1][2] = (m[0][2] * m[1][0] - m[0][0] * m[1][2]) * id;
    inv.m[2][2] = (m[0][0] * m[1][1] - m[0][1] * m[1][0]) * id;
    return true;
}

// cv::Rodrigues, vector -> matrix
M3 rodrigues_to_matrix(const double r[3])
{
    M3 R;
    const double theta = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    if (theta < DBL_EPSILON) {
        std::memset(&R, 0, sizeof R);
        R.m[0][0] = R.m[1][1] = R.m[2][2] = 1.0;
        return R;
    }
    const double c = std::cos(theta), s = std::sin(theta), c1 = 1.0 - c, it = 1.0 / theta;
    const double x = r[0] * it, y = r[1] * it, z = r[2] * it;
    const double rrt[3][3] = {{x * x, x * y, x * z}, {x * y, y * y, y * z}, {x * z, y * z, z * z}};
    const double rx[3][3] = {{0, -z, y}, {z, 0, -x}, {-y, x, 0}};
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) R.m[i][j] = c * (i == j ? 1.0 : 0.0) + c1 * rrt[i][j] + s * rx[i][j];
    return R;
}

// cv::Rodrigues, matrix -> vector.  OpenCV first replaces R by the nearest rotation (U * Vt of its SVD); the same matrix is
// the limit of the Newton iteration X <- (X + X^-T) / 2, which needs no SVD and converges quadratically from a matrix
// that is orthonormal to the digits a YAML file keeps.
bool rodrigues_to_vector(const M3& Rin, double r[3])
{
    M3 R = Rin;
    for (int it = 0; it < 20; it++) {
        M3 inv;
        if (!inverse(R, inv)) return false;
        const M3 invT = transpose(inv);
        double d = 0.0;
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                const double v = 0.5 * (R.m[i][j] + invT.m[i][j]);
                d = std::fmax(d, std::fabs(v - R.m[i][j]));
                R.m[i][j] = v;
            }
        if (d < 1e-17) break;
    }
    double rx = R.m[2][1] - R.m[1][2], ry = R.m[0][2] - R.m[2][0], rz = R.m[1][0] - R.m[0][1];
    const double s = std::sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
    double c = (R.m[0][0] + R.m[1][1] + R.m[2][2] - 1.0) * 0.5;
    c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
    const double theta = std::acos(c);
    if (s < 1e-5) {
        if (c > 0) { r[0] = r[1] = r[2] = 0.0; return true; }
        // rotation by pi: the axis comes from the diagonal
        double t;
        t = (R.m[0][0] + 1) * 0.5; rx = std::sqrt(std::fmax(t, 0.0));
        t = (R.m[1][1] + 1) * 0.5; ry = std::sqrt(std::fmax(t, 0.0)) * (R.m[0][1] < 0 ? -1.0 : 1.0);
        t = (R.m[2][2] + 1) * 0.5; rz = std::sqrt(std::fmax(t, 0.0)) * (R.m[0][2] < 0 ? -1.0 : 1.0);
        if (std::fabs(rx) < std::fabs(ry) && std::fabs(rx) < std::fabs(rz) && (R.m[1][2] > 0) != (ry * rz > 0)) rz = -rz;
        const double k = theta / std::sqrt(rx * rx + ry * ry + rz * rz);
        r[0] = rx * k; r[1] = ry * k; r[2] = rz * k;
        return true;
    }
    const double vth = 1.0 / (2.0 * s) * theta;
    r[0] = rx * vth; r[1] = ry * vth; r[2] = rz * vth;
    return true;
}

// cv::undistortPoints for one point: five fixed-point iterations of the inverse distortion model (default criteria),
// then [x y 1] -> RR [x y 1] with the perspective divide.  k: 14 coefficients, zero padded.
void undistort_point(double u, double v, const double K[9], const double k[14], const M3* RR, double& ox, double& oy)
{
    const double fx = K[0], fy = K[4], cx = K[2], cy = K[5], ifx = 1.0 / fx, ify = 1.0 / fy;
    double x = (u - cx) * ifx, y = (v - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; j++) {
        const double r2 = x * x + y * y;
        const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
        if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
        const double dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
        const double dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
        x = (x0 - dx) * icdist;
        y = (y0 - dy) * icdist;
    }
    if (RR) {
        const double xx = RR->m[0][0] * x + RR->m[0][1] * y + RR->m[0][2];
        const double yy = RR->m[1][0] * x + RR->m[1][1] * y + RR->m[1][2];
        const double ww = 1.0 / (RR->m[2][0] * x + RR->m[2][1] * y + RR->m[2][2]);
        x = xx * ww;
        y = yy * ww;
    }
    ox = x;
    oy = y;
}

struct RectD { double x, y, w, h; };

// Where the image border lands after rectification: 9x9 grid over [0, w-1] x [0, h-1], the largest inscribed and the
// smallest enclosing axis-aligned rectangle of its image.
void undistort_rectangles(const double K[9], const double k[14], const M3& R, const double P[12], int w, int h, RectD& inner, RectD& outer)
{
    const int N = 9;
    M3 PP;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) PP.m[i][j] = P[i * 4 + j];
    const M3 RR = mul(PP, R);
    double iX0 = -FLT_MAX, iX1 = FLT_MAX, iY0 = -FLT_MAX, iY1 = FLT_MAX;
    double oX0 = FLT_MAX, oX1 = -FLT_MAX, oY0 = FLT_MAX, oY1 = -FLT_MAX;
    for (int y = 0; y < N; y++)
        for (int x = 0; x < N; x++) {
            double px, py;
            undistort_point((double)x * (w - 1) / (N - 1), (double)y * (h - 1) / (N - 1), K, k, &RR, px, py);
            oX0 = std::fmin(oX0, px); oX1 = std::fmax(oX1, px);
            oY0 = std::fmin(oY0, py); oY1 = std::fmax(oY1, py);
            if (x == 0) iX0 = std::fmax(iX0, px);
            if (x == N - 1) iX1 = std::fmin(iX1, px);
            if (y == 0) iY0 = std::fmax(iY0, py);
            if (y == N - 1) iY1 = std::fmin(iY1, py);
        }
    inner = {iX0, iY0, iX1 - iX0, iY1 - iY0};
    outer = {oX0, oY0, oX1 - oX0, oY1 - oY0};
}

double max4(double a, double b, double c, double d) { return std::fmax(std::fmax(std::fmax(a, b), c), d); }
double min4(double a, double b, double c, double d) { return std::fmin(std::fmin(std::fmin(a, b), c), d); }

void valid_roi(const RectD& in, double cx0, double cy0, double cx, double cy, double s, int W, int H, int roi[4])
{
    const int rx = (int)std::ceil((in.x - cx0) * s + cx), ry = (int)std::ceil((in.y - cy0) * s + cy);
    const int rw = (int)std::floor(in.w * s), rh = (int)std::floor(in.h * s);
    const int x0 = rx > 0 ? rx : 0, y0 = ry > 0 ? ry : 0;
    const int x1 = rx + rw < W ? rx + rw : W, y1 = ry + rh < H ? ry + rh : H;
    if (x1 <= x0 || y1 <= y0) { roi[0] = roi[1] = roi[2] = roi[3] = 0; return; }       // cv::Rect & cv::Rect: empty -> all zero
    roi[0] = x0; roi[1] = y0; roi[2] = x1 - x0; roi[3] = y1 - y0;
}

}  // namespace

extern "C" {

int sa_stereo_rectify(const double* K_l, const double* D_l, int nD_l, const double* K_r, const double* D_r, int nD_r, int width, int height,
                      const double* R, const double* T, int flags, double alpha, int new_width, int new_height, double* R_L, double* R_R,
                      double* P1, double* P2, double* Q, int* roi_l, int* roi_r)
{
    if (!K_l || !K_r || !R || !T || !R_L || !R_R || !P1 || !P2 || width <= 0 || height <= 0 || nD_l < 0 || nD_r < 0 || nD_l > 14 || nD_r > 14 ||
        (nD_l > 0 && !D_l) || (nD_r > 0 && !D_r) || new_width < 0 || new_height < 0)
        return SA_ERR_INVALID_ARG;
    double k[2][14] = {{0}, {0}};
    for (int i = 0; i < nD_l; i++) k[0][i] = D_l[i];
    for (int i = 0; i < nD_r; i++) k[1][i] = D_r[i];
    const double* K[2] = {K_l, K_r};
    const double nx = width, ny = height;

    // both cameras are turned half of R towards each other ...
    M3 Rm;
    std::memcpy(Rm.m, R, sizeof Rm.m);
    double om[3];
    if (!rodrigues_to_vector(Rm, om)) return SA_ERR_INVALID_ARG;
    for (int i = 0; i < 3; i++) om[i] *= -0.5;
    const M3 r_r = rodrigues_to_matrix(om);
    double t[3];
    mul_vec(r_r, T, t);
    // ... then about the axis that brings the baseline onto the x (or y) axis
    const int idx = std::fabs(t[0]) > std::fabs(t[1]) ? 0 : 1;
    const double c = t[idx], nt = std::sqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
    if (!(nt > 0.0)) return SA_ERR_INVALID_ARG;
    double uu[3] = {0, 0, 0};
    uu[idx] = c > 0 ? 1 : -1;
    double ww[3] = {t[1] * uu[2] - t[2] * uu[1], t[2] * uu[0] - t[0] * uu[2], t[0] * uu[1] - t[1] * uu[0]};
    const double nw = std::sqrt(ww[0] * ww[0] + ww[1] * ww[1] + ww[2] * ww[2]);
    if (nw > 0.0) {
        const double sc = std::acos(std::fabs(c) / nt) / nw;
        for (int i = 0; i < 3; i++) ww[i] *= sc;
    }
    const M3 wR = rodrigues_to_matrix(ww);
    const M3 R1 = mul(wR, transpose(r_r)), R2 = mul(wR, r_r);
    mul_vec(R2, T, t);

    // new focal length and principal points
    if ((long long)new_width * new_height == 0) { new_width = width; new_height = height; }
    const double ratio_x = (double)new_width / width / 2, ratio_y = (double)new_height / height / 2;
    const double ratio = idx == 1 ? ratio_x : ratio_y;
    double fc_new = (K_l[(idx ^ 1) * 4] + K_r[(idx ^ 1) * 4]) * ratio;
    double cc[2][2];
    for (int e = 0; e < 2; e++) {
        const M3& Re = e == 0 ? R1 : R2;
        double ax = 0.0, ay = 0.0;
        for (int i = 0; i < 4; i++) {
            // the four image corners, undistorted; OpenCV keeps them (and their projections) as float
            const float u = (float)((i % 2) * (nx - 1)), v = (float)((i < 2 ? 0 : 1) * (ny - 1));
            double x, y;
            undistort_point(u, v, K[e], k[e], nullptr, x, y);
            const double X[3] = {(double)(float)x, (double)(float)y, 1.0};
            double Y[3];
            mul_vec(Re, X, Y);
            const double z = Y[2] != 0.0 ? 1.0 / Y[2] : 1.0;
            ax += (double)(float)(fc_new * (Y[0] * z));
            ay += (double)(float)(fc_new * (Y[1] * z));
        }
        cc[e][0] = (nx - 1) / 2 - ax / 4;
        cc[e][1] = (ny - 1) / 2 - ay / 4;
    }
    if (flags & SA_CALIB_ZERO_DISPARITY) {
        cc[0][0] = cc[1][0] = (cc[0][0] + cc[1][0]) * 0.5;
        cc[0][1] = cc[1][1] = (cc[0][1] + cc[1][1]) * 0.5;
    } else if (idx == 0) {
        cc[0][1] = cc[1][1] = (cc[0][1] + cc[1][1]) * 0.5;        // horizontal stereo
    } else {
        cc[0][0] = cc[1][0] = (cc[0][0] + cc[1][0]) * 0.5;        // vertical stereo
    }
    double p1[12] = {fc_new, 0, cc[0][0], 0, 0, fc_new, cc[0][1], 0, 0, 0, 1, 0};
    double p2[12] = {fc_new, 0, cc[1][0], 0, 0, fc_new, cc[1][1], 0, 0, 0, 1, 0};
    p2[idx * 4 + 3] = t[idx] * fc_new;                             // baseline * focal length

    // alpha: scale so that only valid pixels remain (0), all source pixels are kept (1), or leave the scale alone (< 0)
    if (alpha > 1.0) alpha = 1.0;
    RectD in1, out1, in2, out2;
    undistort_rectangles(K_l, k[0], R1, p1, width, height, in1, out1);
    undistort_rectangles(K_r, k[1], R2, p2, width, height, in2, out2);
    const double cx1_0 = cc[0][0], cy1_0 = cc[0][1], cx2_0 = cc[1][0], cy2_0 = cc[1][1];
    const double cx1 = new_width * cx1_0 / width, cy1 = new_height * cy1_0 / height;
    const double cx2 = new_width * cx2_0 / width, cy2 = new_height * cy2_0 / height;
    double s = 1.0;
    if (alpha >= 0) {
        const int W1 = new_width - 1, H1 = new_height - 1;
        double s0 = max4(cx1 / (cx1_0 - in1.x), cy1 / (cy1_0 - in1.y), (W1 - cx1) / (in1.x + in1.w - cx1_0), (H1 - cy1) / (in1.y + in1.h - cy1_0));
        s0 = std::fmax(max4(cx2 / (cx2_0 - in2.x), cy2 / (cy2_0 - in2.y), (W1 - cx2) / (in2.x + in2.w - cx2_0), (H1 - cy2) / (in2.y + in2.h - cy2_0)), s0);
        double s1 = min4(cx1 / (cx1_0 - out1.x), cy1 / (cy1_0 - out1.y), (W1 - cx1) / (out1.x + out1.w - cx1_0), (H1 - cy1) / (out1.y + out1.h - cy1_0));
        s1 = std::fmin(min4(cx2 / (cx2_0 - out2.x), cy2 / (cy2_0 - out2.y), (W1 - cx2) / (out2.x + out2.w - cx2_0), (H1 - cy2) / (out2.y + out2.h - cy2_0)), s1);
        s = s0 * (1 - alpha) + s1 * alpha;
    }
    fc_new *= s;
    p1[0] = p1[5] = fc_new; p1[2] = cx1; p1[6] = cy1;
    p2[0] = p2[5] = fc_new; p2[2] = cx2; p2[6] = cy2;
    p2[idx * 4 + 3] = s * p2[idx * 4 + 3];

    std::memcpy(R_L, R1.m, sizeof R1.m);
    std::memcpy(R_R, R2.m, sizeof R2.m);
    std::memcpy(P1, p1, sizeof p1);
    std::memcpy(P2, p2, sizeof p2);
    if (roi_l) valid_roi(in1, cx1_0, cy1_0, cx1, cy1, s, new_width, new_height, roi_l);
    if (roi_r) valid_roi(in2, cx2_0, cy2_0, cx2, cy2, s, new_width, new_height, roi_r);
    if (Q) {
        const double q[16] = {1, 0, 0, -cx1, 0, 1, 0, -cy1, 0, 0, 0, fc_new, 0, 0, -1.0 / t[idx], (idx == 0 ? cx1 - cx2 : cy1 - cy2) / t[idx]};
        std::memcpy(Q, q, sizeof q);
    }
    return SA_OK;
}

int sa_rectify_calibration(sa_calibration* c, int width, int height, int flags, double alpha, int new_width, int new_height, int* roi_l, int* roi_r)
{
    if (!c) return SA_ERR_INVALID_ARG;
    return sa_stereo_rectify(c->K_l, c->D_l, c->nD_l, c->K_r, c->D_r, c->nD_r, width, height, c->R, c->T, flags, alpha, new_width, new_height, c->R_L, c->R_R,
                             c->P1, c->P2, c->Q, roi_l, roi_r);
}

}  // extern "C"
