/*
 * TEST INFRASTRUCTURE ONLY (oracle build). Not part of the product path.
 *
 * Software stand-in for the seven OpenGL 1.1 fixed-function matrix entry points that the
 * reference's ViewBase::matrix_rotate uses purely as a 4x4 matrix library
 * (reference ViewBase.cpp:34-47: glMatrixMode, glPushMatrix, glPopMatrix, glLoadIdentity,
 * glLoadMatrixf, glGetFloatv, glRotatef). This is matrix arithmetic, not a graphics API: no
 * context, no loader, no device. It exists so the reference's own ViewBase.cpp links headless.
 *
 * Semantics implemented (OpenGL 1.1 spec, section 2.10.2): matrices are column-major float[16];
 * glRotatef(a, x, y, z) post-multiplies the current matrix by the rotation of `a` degrees about
 * the normalised axis (x, y, z):  M <- M * R.
 *
 * The fp32 rounding of this shim is NOT parity relevant: the oracle and the CUDA kernel both
 * consume the resulting `View` struct bit-for-bit.
 */
#include <math.h>
#include <string.h>

#define SHIM_STACK_DEPTH 32

static float g_stack[SHIM_STACK_DEPTH][16] = {
    {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}};
static int g_top = 0;

void glMatrixMode(unsigned int mode) { (void)mode; }

void glPushMatrix(void) {
    if (g_top + 1 < SHIM_STACK_DEPTH) {
        memcpy(g_stack[g_top + 1], g_stack[g_top], sizeof(g_stack[0]));
        g_top++;
    }
}

void glPopMatrix(void) {
    if (g_top > 0) g_top--;
}

void glLoadIdentity(void) {
    static const float id[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    memcpy(g_stack[g_top], id, sizeof(id));
}

void glLoadMatrixf(const float *m) { memcpy(g_stack[g_top], m, 16 * sizeof(float)); }

void glGetFloatv(unsigned int pname, float *out) {
    (void)pname; /* only GL_MODELVIEW_MATRIX is ever requested (ViewBase.cpp:39,45) */
    memcpy(out, g_stack[g_top], 16 * sizeof(float));
}

void glRotatef(float angle, float x, float y, float z) {
    float len = sqrtf(x * x + y * y + z * z);
    if (len == 0.0f) return; /* degenerate axis: GL leaves the matrix unchanged in practice */
    x /= len; y /= len; z /= len;
    float rad = angle * 3.14159265358979323846f / 180.0f;
    float c = cosf(rad), s = sinf(rad), t = 1.0f - c;
    /* column-major rotation matrix */
    float r[16] = {
        x * x * t + c,     y * x * t + z * s, x * z * t - y * s, 0,
        x * y * t - z * s, y * y * t + c,     y * z * t + x * s, 0,
        x * z * t + y * s, y * z * t - x * s, z * z * t + c,     0,
        0, 0, 0, 1};
    float *m = g_stack[g_top];
    float o[16];
    for (int col = 0; col < 4; col++)
        for (int row = 0; row < 4; row++) {
            float acc = 0;
            for (int k = 0; k < 4; k++) acc += m[k * 4 + row] * r[col * 4 + k];
            o[col * 4 + row] = acc;
        }
    memcpy(m, o, sizeof(o));
}
