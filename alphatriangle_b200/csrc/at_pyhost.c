/* Host side of the evaluate_batch boundary in C (CPython extension `alphatriangle_b200._pyhost`).
 *
 * Two loops that bound the API-level throughput once the GPU path takes 0.16 us per state:
 *
 *   pack_states(states, rows, cols, slots, out)
 *       list[GameState] -> PackedState records (alphatriangle_b200/packing.py holds the format and the Python twin of
 *       this loop). Reads exactly the accessor surface the reference's extractor reads
 *       (alphatriangle/features/extractor.py:28-31, 58-66, 109, 114): get_grid_data_np()["occupied" / "death"],
 *       get_shapes(), shape.triangles, shape.bbox(), game_score(), current_step.
 *
 *   build_results(probs, values, n, actions, keys)
 *       float32 [n, actions] + float32 [n] -> list[(dict[int, float], float)] with max(p, 0) per entry
 *       (alphatriangle/nn/network.py:281-311), plus the indices of the rows whose clipped sum is further than 5e-6
 *       from 1 -- those the caller sends through the reference's exact per-row rule (renormalise / uniform fallback).
 *
 * No CUDA here; the module is pure host code and is built next to libat_b200.so by the same Makefile.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#define AT_STATE_BYTES 80
#define AT_MAX_CELLS 128
#define AT_MAX_SLOTS 4

static PyObject *s_get_grid, *s_get_shapes, *s_triangles, *s_bbox, *s_game_score, *s_current_step, *s_occupied, *s_death;

/* bool-like [rows, cols] array (numpy bool / uint8 / int8, any strides) -> 16-byte little-endian bit mask.
 * Returns 0 on success, -1 with a Python error set, 1 if the object is not a plain byte array (caller falls back). */
static int mask_bits(PyObject* arr, int rows, int cols, uint8_t out[16], const char* what) {
  Py_buffer v;
  if (PyObject_GetBuffer(arr, &v, PyBUF_STRIDES | PyBUF_FORMAT) != 0) {
    PyErr_Clear();
    return 1;
  }
  int rc = 0;
  if (v.itemsize != 1 || (v.format && v.format[0] != '?' && v.format[0] != 'B' && v.format[0] != 'b')) {
    rc = 1;
  } else if (v.ndim != 2 || v.shape[0] != rows || v.shape[1] != cols) {
    rc = 1; /* the Python packer words the shape error; keep one source of messages */
  } else {
    memset(out, 0, 16);
    const char* base = (const char*)v.buf;
    for (int r = 0; r < rows; ++r)
      for (int c = 0; c < cols; ++c) {
        const int i = r * cols + c;
        if (base[r * v.strides[0] + c * v.strides[1]] != 0) out[i >> 3] |= (uint8_t)(1u << (i & 7));
      }
  }
  PyBuffer_Release(&v);
  (void)what;
  return rc;
}

static int as_i8(PyObject* o, int8_t* dst) {
  const long x = PyLong_AsLong(o);
  if (x == -1 && PyErr_Occurred()) return -1;
  if (x < -128 || x > 127) {
    PyErr_Format(PyExc_ValueError, "shape bbox coordinate %ld outside int8", x);
    return -1;
  }
  *dst = (int8_t)x;
  return 0;
}

/* one state -> one record; 0 ok, -1 error set, 1 unsupported object layout (caller falls back to the Python packer) */
static int pack_one(PyObject* state, int rows, int cols, int slots, uint8_t* rec) {
  memset(rec, 0, AT_STATE_BYTES);
  PyObject* grid = PyObject_CallMethodNoArgs(state, s_get_grid);
  if (!grid) return -1;
  PyObject* occ = PyObject_GetItem(grid, s_occupied);
  PyObject* death = occ ? PyObject_GetItem(grid, s_death) : NULL;
  Py_DECREF(grid);
  if (!occ || !death) {
    Py_XDECREF(occ);
    Py_XDECREF(death);
    return -1;
  }
  int rc = mask_bits(occ, rows, cols, rec, "occupied");
  if (rc == 0) rc = mask_bits(death, rows, cols, rec + 16, "death");
  Py_DECREF(occ);
  Py_DECREF(death);
  if (rc != 0) return rc;

  PyObject* shapes = PyObject_CallMethodNoArgs(state, s_get_shapes);
  if (!shapes) return -1;
  PyObject* seq = PySequence_Fast(shapes, "get_shapes() must return a sequence");
  Py_DECREF(shapes);
  if (!seq) return -1;
  const Py_ssize_t n_shapes = PySequence_Fast_GET_SIZE(seq);
  if (n_shapes > slots) {
    PyErr_Format(PyExc_ValueError, "state has %zd shape slots, EnvConfig says %d", n_shapes, slots);
    Py_DECREF(seq);
    return -1;
  }
  for (Py_ssize_t i = 0; i < n_shapes; ++i) {
    PyObject* shape = PySequence_Fast_GET_ITEM(seq, i); /* borrowed */
    const int truth = PyObject_IsTrue(shape);
    if (truth < 0) goto fail_seq;
    if (!truth) continue;
    int8_t* s = (int8_t*)(rec + 32 + 8 * i);
    s[0] = 1;
    PyObject* tris = PyObject_GetAttr(shape, s_triangles);
    if (!tris) goto fail_seq;
    PyObject* tseq = PySequence_Fast(tris, "shape.triangles must be a sequence");
    Py_DECREF(tris);
    if (!tseq) goto fail_seq;
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(tseq);
    if (n == 0) {
      Py_DECREF(tseq);
      continue;
    }
    if (n > 255) {
      PyErr_Format(PyExc_ValueError, "shape in slot %zd has %zd triangles; PackedState holds at most 255", i, n);
      Py_DECREF(tseq);
      goto fail_seq;
    }
    int ups = 0;
    for (Py_ssize_t t = 0; t < n; ++t) {
      PyObject* tri = PySequence_Fast_GET_ITEM(tseq, t); /* (r, c, is_up) */
      PyObject* up = PySequence_GetItem(tri, 2);
      if (!up) {
        Py_DECREF(tseq);
        goto fail_seq;
      }
      const int u = PyObject_IsTrue(up);
      Py_DECREF(up);
      if (u < 0) {
        Py_DECREF(tseq);
        goto fail_seq;
      }
      ups += u;
    }
    Py_DECREF(tseq);
    PyObject* bbox = PyObject_CallMethodNoArgs(shape, s_bbox);
    if (!bbox) goto fail_seq;
    PyObject* bseq = PySequence_Fast(bbox, "shape.bbox() must return four integers");
    Py_DECREF(bbox);
    if (!bseq) goto fail_seq;
    if (PySequence_Fast_GET_SIZE(bseq) != 4) {
      PyErr_SetString(PyExc_ValueError, "shape.bbox() must return four integers");
      Py_DECREF(bseq);
      goto fail_seq;
    }
    for (int k = 0; k < 4; ++k)
      if (as_i8(PySequence_Fast_GET_ITEM(bseq, k), &s[4 + k]) != 0) {
        Py_DECREF(bseq);
        goto fail_seq;
      }
    Py_DECREF(bseq);
    s[1] = (int8_t)(uint8_t)n;
    s[2] = (int8_t)(uint8_t)ups;
  }
  Py_DECREF(seq);

  {
    PyObject* sc = PyObject_CallMethodNoArgs(state, s_game_score);
    if (!sc) return -1;
    const double score = PyFloat_AsDouble(sc);
    Py_DECREF(sc);
    if (score == -1.0 && PyErr_Occurred()) return -1;
    memcpy(rec + 64, &score, 8);
    PyObject* st = PyObject_GetAttr(state, s_current_step);
    if (!st) return -1;
    PyObject* st_int = PyNumber_Long(st);
    Py_DECREF(st);
    if (!st_int) return -1;
    int overflow = 0;
    long long step = PyLong_AsLongLongAndOverflow(st_int, &overflow);
    Py_DECREF(st_int);
    if (step == -1 && !overflow && PyErr_Occurred()) return -1;
    if (overflow > 0 || step > INT32_MAX) step = INT32_MAX;
    if (overflow < 0 || step < INT32_MIN) step = INT32_MIN;
    const int32_t s32 = (int32_t)step;
    memcpy(rec + 72, &s32, 4);
  }
  return 0;

fail_seq:
  Py_DECREF(seq);
  return -1;
}

/* pack_states(states, rows, cols, slots, out) -> number of states packed, or -1 - index of the first state whose arrays
 * are not plain byte arrays of the expected shape (the Python packer then handles the batch and words the error). */
static PyObject* py_pack_states(PyObject* self, PyObject* args) {
  PyObject *states, *out;
  int rows, cols, slots;
  if (!PyArg_ParseTuple(args, "OiiiO", &states, &rows, &cols, &slots, &out)) return NULL;
  if (rows <= 0 || cols <= 0 || rows * cols > AT_MAX_CELLS || slots <= 0 || slots > AT_MAX_SLOTS) {
    PyErr_SetString(PyExc_ValueError, "pack_states: board or slot count outside the PackedState limits");
    return NULL;
  }
  PyObject* seq = PySequence_Fast(states, "states must be a sequence");
  if (!seq) return NULL;
  const Py_ssize_t n = PySequence_Fast_GET_SIZE(seq);
  Py_buffer ov;
  if (PyObject_GetBuffer(out, &ov, PyBUF_WRITABLE | PyBUF_C_CONTIGUOUS) != 0) {
    Py_DECREF(seq);
    return NULL;
  }
  PyObject* result = NULL;
  if (ov.len != n * AT_STATE_BYTES) {
    PyErr_Format(PyExc_ValueError, "out holds %zd bytes, %zd states need %zd", ov.len, n, n * (Py_ssize_t)AT_STATE_BYTES);
  } else {
    Py_ssize_t i = 0;
    int rc = 0;
    for (; i < n; ++i) {
      rc = pack_one(PySequence_Fast_GET_ITEM(seq, i), rows, cols, slots, (uint8_t*)ov.buf + i * AT_STATE_BYTES);
      if (rc != 0) break;
    }
    if (rc == 0) result = PyLong_FromSsize_t(n);
    else if (rc > 0) result = PyLong_FromSsize_t(-1 - i);
  }
  PyBuffer_Release(&ov);
  Py_DECREF(seq);
  return result;
}

/* build_results(probs, values, n, actions, keys) -> (list[(dict, float)], list[int] rows to re-check) */
static PyObject* py_build_results(PyObject* self, PyObject* args) {
  PyObject *probs_o, *values_o, *keys;
  Py_ssize_t n, actions;
  if (!PyArg_ParseTuple(args, "OOnnO!", &probs_o, &values_o, &n, &actions, &PyTuple_Type, &keys)) return NULL;
  if (n < 0 || actions <= 0 || PyTuple_GET_SIZE(keys) != actions) {
    PyErr_SetString(PyExc_ValueError, "build_results: keys must be a tuple with one entry per action");
    return NULL;
  }
  Py_buffer pv, vv;
  if (PyObject_GetBuffer(probs_o, &pv, PyBUF_C_CONTIGUOUS | PyBUF_FORMAT) != 0) return NULL;
  if (PyObject_GetBuffer(values_o, &vv, PyBUF_C_CONTIGUOUS | PyBUF_FORMAT) != 0) {
    PyBuffer_Release(&pv);
    return NULL;
  }
  PyObject *results = NULL, *recheck = NULL, *ret = NULL;
  if (pv.itemsize != 4 || vv.itemsize != 4 || !pv.format || !vv.format || pv.format[strlen(pv.format) - 1] != 'f' ||
      vv.format[strlen(vv.format) - 1] != 'f' || pv.len != n * actions * 4 || vv.len != n * 4) {
    PyErr_SetString(PyExc_ValueError, "build_results: probs must be float32 [n, actions] and values float32 [n], C-contiguous");
    goto done;
  }
  results = PyList_New(n);
  recheck = PyList_New(0);
  if (!results || !recheck) goto done;
  {
    const float* P = (const float*)pv.buf;
    const float* V = (const float*)vv.buf;
    for (Py_ssize_t i = 0; i < n; ++i) {
      const float* row = P + i * actions;
      PyObject* d = _PyDict_NewPresized(actions);
      if (!d) goto done;
      float sum = 0.0f; /* float32 accumulation like numpy's sum over a float32 row is not bit-identical, hence the loose 5e-6 gate */
      int bad = 0;
      for (Py_ssize_t a = 0; a < actions; ++a) {
        float p = row[a];
        if (!(p > 0.0f)) { /* max(p, 0); NaN also lands here and sends the row to the exact rule */
          if (p != p) bad = 1;
          p = 0.0f;
        }
        sum += p;
        PyObject* f = PyFloat_FromDouble((double)p);
        if (!f || PyDict_SetItem(d, PyTuple_GET_ITEM(keys, a), f) != 0) {
          Py_XDECREF(f);
          Py_DECREF(d);
          goto done;
        }
        Py_DECREF(f);
      }
      if (bad || !(fabsf(sum - 1.0f) <= 5e-6f)) {
        PyObject* idx = PyLong_FromSsize_t(i);
        if (!idx || PyList_Append(recheck, idx) != 0) {
          Py_XDECREF(idx);
          Py_DECREF(d);
          goto done;
        }
        Py_DECREF(idx);
      }
      PyObject* val = PyFloat_FromDouble((double)V[i]);
      PyObject* tup = val ? PyTuple_Pack(2, d, val) : NULL;
      Py_XDECREF(val);
      Py_DECREF(d);
      if (!tup) goto done;
      PyList_SET_ITEM(results, i, tup);
    }
  }
  ret = PyTuple_Pack(2, results, recheck);
done:
  Py_XDECREF(results);
  Py_XDECREF(recheck);
  PyBuffer_Release(&pv);
  PyBuffer_Release(&vv);
  return ret;
}

static PyMethodDef methods[] = {
    {"pack_states", py_pack_states, METH_VARARGS, "pack_states(states, rows, cols, slots, out) -> n, or -1 - i for an unsupported state i"},
    {"build_results", py_build_results, METH_VARARGS, "build_results(probs, values, n, actions, keys) -> (results, rows to re-check)"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef module = {PyModuleDef_HEAD_INIT, "_pyhost", "C host loops of the evaluate_batch boundary", -1, methods};

PyMODINIT_FUNC PyInit__pyhost(void) {
  s_get_grid = PyUnicode_InternFromString("get_grid_data_np");
  s_get_shapes = PyUnicode_InternFromString("get_shapes");
  s_triangles = PyUnicode_InternFromString("triangles");
  s_bbox = PyUnicode_InternFromString("bbox");
  s_game_score = PyUnicode_InternFromString("game_score");
  s_current_step = PyUnicode_InternFromString("current_step");
  s_occupied = PyUnicode_InternFromString("occupied");
  s_death = PyUnicode_InternFromString("death");
  if (!s_get_grid || !s_get_shapes || !s_triangles || !s_bbox || !s_game_score || !s_current_step || !s_occupied || !s_death) return NULL;
  return PyModule_Create(&module);
}
