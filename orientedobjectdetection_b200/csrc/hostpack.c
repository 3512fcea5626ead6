/* _hostpack: host side of the huge-image patch merge (reference
 * mmrotate/core/patch/merge_results.py:7-29,98-107): the reference walks patches x classes in
 * Python, shifts every (k,6) array by its patch origin and concatenates per class — 6000 small
 * numpy calls for a 400-patch x 15-class scene.  gather_shift() does the same walk in C through
 * the buffer protocol: one pass validates every array (nothing is touched if any is not a
 * writable 2-D float32 (k,6) array), the second shifts cx, cy IN PLACE (the caller's arrays
 * change exactly as the reference's translate_bboxes changes them) and copies the rows, class
 * after class, patch after patch inside a class, into one caller-owned staging buffer (pinned
 * host memory of the upload).
 *
 * Arithmetic of the shift = numpy's for float32 rows + an int64 / float64 offset:
 * (float)((double)x + offset).
 *
 * Plain CPython C API, no numpy headers.  Host code only: nothing here computes geometry.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define HP_COLS 6

static int cmp_ptr(const void *a, const void *b) {
    const uintptr_t x = *(const uintptr_t *)a, y = *(const uintptr_t *)b;
    return x < y ? -1 : x > y;
}

static int is_f32(const Py_buffer *v) {
    const char *f = v->format;
    if (v->itemsize != 4) return 0;
    if (!f) return 0;
    if (*f == '@' || *f == '=' || *f == '<') ++f;   /* native / little endian (the image is x86-64) */
    return f[0] == 'f' && f[1] == 0;
}

/* gather_shift(results, offsets, out[, take]) -> (total_rows, counts, filled) | None
 *   results  sequence over patches of sequences over classes of (k,6) float32 arrays
 *   offsets  contiguous float64 buffer, 2 values per patch
 *   out      writable contiguous buffer (the float32 staging rows)
 *   take     optional bytes-like, one byte per class: rows of classes with a zero byte are shifted
 *            in place like all others but not copied (multi-GPU: a rank stages its own classes);
 *            total_rows then counts the copied classes only, counts still lists every class
 * None: some array is not a writable 2-D float32 array with 6 columns, or the same non-empty array appears
 * twice (caller takes the numpy path); filled False: `out` is smaller than total_rows * 24 bytes — nothing was shifted. */
static PyObject *gather_shift(PyObject *self, PyObject *args) {
    PyObject *results, *offs_obj, *out_obj, *take_obj = NULL;
    if (!PyArg_ParseTuple(args, "OOO|O", &results, &offs_obj, &out_obj, &take_obj)) return NULL;
    if (take_obj == Py_None) take_obj = NULL;
    Py_buffer offs, out, take;
    if (PyObject_GetBuffer(offs_obj, &offs, PyBUF_C_CONTIGUOUS | PyBUF_FORMAT) < 0) return NULL;
    if (PyObject_GetBuffer(out_obj, &out, PyBUF_C_CONTIGUOUS | PyBUF_WRITABLE) < 0) {
        PyBuffer_Release(&offs);
        return NULL;
    }
    if (take_obj && PyObject_GetBuffer(take_obj, &take, PyBUF_C_CONTIGUOUS) < 0) {
        PyBuffer_Release(&out);
        PyBuffer_Release(&offs);
        return NULL;
    }
    const unsigned char *takeb = take_obj ? (const unsigned char *)take.buf : NULL;
    PyObject *res_fast = PySequence_Fast(results, "results must be a sequence over patches");
    PyObject *ret = NULL, *counts = NULL;
    PyObject **rows_fast = NULL;
    Py_buffer *views = NULL;
    Py_ssize_t nviews = 0, P = 0, C = 0;
    int unsupported = 0;
    if (!res_fast) goto done;
    P = PySequence_Fast_GET_SIZE(res_fast);
    if (offs.itemsize != 8 || !offs.format || strcmp(offs.format, "d") != 0 || offs.len != P * 16) {
        PyErr_SetString(PyExc_ValueError, "offsets must be a contiguous float64 buffer with 2 values per patch");
        goto done;
    }
    if (P == 0) {
        ret = Py_BuildValue("(n[]O)", (Py_ssize_t)0, Py_True);
        goto done;
    }
    rows_fast = (PyObject **)calloc((size_t)P, sizeof(PyObject *));
    if (!rows_fast) { PyErr_NoMemory(); goto done; }
    for (Py_ssize_t i = 0; i < P; ++i) {
        rows_fast[i] = PySequence_Fast(PySequence_Fast_GET_ITEM(res_fast, i), "a patch result must be a sequence over classes");
        if (!rows_fast[i]) goto done;
        const Py_ssize_t c = PySequence_Fast_GET_SIZE(rows_fast[i]);
        if (i == 0) C = c;
        else if (c != C) {
            PyErr_SetString(PyExc_ValueError, "every patch must hold the same number of classes");
            goto done;
        }
    }
    if (takeb && take.len != C) {
        PyErr_SetString(PyExc_ValueError, "take must hold one byte per class");
        goto done;
    }
    views = (Py_buffer *)malloc((size_t)(P * C ? P * C : 1) * sizeof(Py_buffer));
    if (!views) { PyErr_NoMemory(); goto done; }
    /* pass 1: views + validation + rows per class; nothing is modified yet */
    int64_t total = 0;
    int64_t *cls_rows = (int64_t *)calloc((size_t)(C ? C : 1), sizeof(int64_t));
    if (!cls_rows) { PyErr_NoMemory(); goto done; }
    for (Py_ssize_t i = 0; i < P && !unsupported; ++i)
        for (Py_ssize_t c = 0; c < C; ++c) {
            Py_buffer *v = &views[i * C + c];
            if (PyObject_GetBuffer(PySequence_Fast_GET_ITEM(rows_fast[i], c), v, PyBUF_RECORDS) < 0) {
                PyErr_Clear();          /* not a buffer / read only: the numpy path raises what the reference raises */
                unsupported = 1;
                break;
            }
            ++nviews;
            if (v->ndim != 2 || v->shape[1] != HP_COLS || !is_f32(v)) {
                unsupported = 1;
                break;
            }
            cls_rows[c] += v->shape[0];
            if (!takeb || takeb[c]) total += v->shape[0];
        }
    if (unsupported) {
        free(cls_rows);
        ret = Py_None;
        Py_INCREF(ret);
        goto done;
    }
    {
        /* the same non-empty array twice in `results`: the reference's loop shifts it twice and concatenates the
         * final state twice (it collects references and concatenates afterwards) — left to the numpy walk */
        const Py_ssize_t items = P * C;
        uintptr_t *ptrs = (uintptr_t *)malloc((size_t)(items ? items : 1) * sizeof(uintptr_t));
        int aliased = 0;
        if (!ptrs) { free(cls_rows); PyErr_NoMemory(); goto done; }
        Py_ssize_t m = 0;
        for (Py_ssize_t t = 0; t < items; ++t)
            if (views[t].shape[0] > 0) ptrs[m++] = (uintptr_t)views[t].buf;
        qsort(ptrs, (size_t)m, sizeof(uintptr_t), cmp_ptr);
        for (Py_ssize_t t = 1; t < m; ++t)
            if (ptrs[t] == ptrs[t - 1]) { aliased = 1; break; }
        free(ptrs);
        if (aliased) {
            free(cls_rows);
            ret = Py_None;
            Py_INCREF(ret);
            goto done;
        }
    }
    counts = PyList_New(C);
    if (!counts) { free(cls_rows); goto done; }
    for (Py_ssize_t c = 0; c < C; ++c) PyList_SET_ITEM(counts, c, PyLong_FromLongLong(cls_rows[c]));
    free(cls_rows);
    if ((int64_t)out.len < total * HP_COLS * 4) {
        ret = Py_BuildValue("(LOO)", (long long)total, counts, Py_False);
        goto done;
    }
    /* pass 2: shift in place + copy, class-major.  Every (class, patch) array has its own destination range, so
     * the 6000 items are independent: a few OpenMP threads share them (memory-bound; the GIL stays with the caller,
     * nothing below touches a Python object). */
    {
        float *dst0 = (float *)out.buf;
        const double *off = (const double *)offs.buf;
        int64_t *start = (int64_t *)malloc((size_t)(P * C ? P * C : 1) * sizeof(int64_t));
        if (!start) { PyErr_NoMemory(); goto done; }
        int64_t run = 0;
        for (Py_ssize_t c = 0; c < C; ++c)
            for (Py_ssize_t i = 0; i < P; ++i) {
                start[c * P + i] = run;
                if (!takeb || takeb[c]) run += views[i * C + c].shape[0];
            }
        const Py_ssize_t items = P * C;
        const int threads = total >= 20000 ? 4 : 1;
        (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 32) num_threads(threads)
#endif
        for (Py_ssize_t t = 0; t < items; ++t) {
            const Py_ssize_t c = t / P, i = t % P;
            const Py_buffer *v = &views[i * C + c];
            const Py_ssize_t k = v->shape[0], s0 = v->strides[0], s1 = v->strides[1];
            const double ox = off[2 * i], oy = off[2 * i + 1];
            char *row = (char *)v->buf;
            float *dst = dst0 + start[t] * HP_COLS;
            if (takeb && !takeb[c]) {           /* somebody else's class: the in-place shift only */
                for (Py_ssize_t r = 0; r < k; ++r, row += s0) {
                    float *px = (float *)row, *py = (float *)(row + s1);
                    *px = (float)((double)*px + ox);
                    *py = (float)((double)*py + oy);
                }
            } else if (s1 == 4) {
                for (Py_ssize_t r = 0; r < k; ++r, row += s0, dst += HP_COLS) {
                    float *p = (float *)row;
                    p[0] = (float)((double)p[0] + ox);
                    p[1] = (float)((double)p[1] + oy);
                    memcpy(dst, p, HP_COLS * sizeof(float));
                }
            } else {
                for (Py_ssize_t r = 0; r < k; ++r, row += s0, dst += HP_COLS) {
                    float *px = (float *)row, *py = (float *)(row + s1);
                    *px = (float)((double)*px + ox);
                    *py = (float)((double)*py + oy);
                    for (int j = 0; j < HP_COLS; ++j) dst[j] = *(float *)(row + j * s1);
                }
            }
        }
        free(start);
    }
    ret = Py_BuildValue("(LOO)", (long long)total, counts, Py_True);
done:
    Py_XDECREF(counts);
    if (views) {
        for (Py_ssize_t j = 0; j < nviews; ++j) PyBuffer_Release(&views[j]);
        free(views);
    }
    if (rows_fast) {
        for (Py_ssize_t i = 0; i < P; ++i) Py_XDECREF(rows_fast[i]);
        free(rows_fast);
    }
    Py_XDECREF(res_fast);
    if (take_obj) PyBuffer_Release(&take);
    PyBuffer_Release(&out);
    PyBuffer_Release(&offs);
    return ret;
}

static PyMethodDef methods[] = {
    {"gather_shift", gather_shift, METH_VARARGS,
     "gather_shift(results, offsets_f64, out[, take]) -> (total_rows, rows_per_class, filled) | None"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_hostpack", "host-side row gathering of the patch merge", -1, methods};

PyMODINIT_FUNC PyInit__hostpack(void) { return PyModule_Create(&moddef); }
