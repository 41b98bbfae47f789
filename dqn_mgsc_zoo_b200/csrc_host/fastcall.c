/* Optional CPython helper for the three per-step calls of the drop-in replay / learner
 * (replay.add -> dqnb_replay_put, learner.update_from_replay -> dqnb_learner_update_from_replay,
 * learner.loss_lagged -> dqnb_learner_loss_post / _take).
 *
 * It calls the SAME C-ABI entry points of libdqnb200.so that the ctypes wrappers call (their addresses are handed
 * over from Python as integers): what it removes is the Python-side cost of one call - ctypes argument
 * conversion, `arr.ctypes.data`, np.ascontiguousarray, int()/float() of NumPy scalars - about 3 us of the 9 us
 * of a replay.add at a 1 M replay.  The e2e step is paced by the host (4 adds + sample + update per 64 us device
 * step), so those microseconds are the headline.  Absent module (not built): the ctypes path is used.
 *
 * Return convention: the C-ABI status (0 = ok), or kFallback when the arguments need the general path
 * (non-contiguous arrays, other dtypes/shapes): the Python caller then runs the ctypes code, which raises the
 * proper error or converts.  No new behaviour lives here. */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>

enum { kFallback = -1000 };

typedef int (*put_fn)(void*, int64_t, const uint8_t*, const uint8_t*, int32_t, float, float, uintptr_t);
typedef int (*update_fn)(void*, void*, const int64_t*, uintptr_t);
typedef int (*post_fn)(void*, uintptr_t);
typedef int (*take_fn)(void*, float*);

static int as_ptr(PyObject* o, void** out) {
  unsigned long long v = PyLong_AsUnsignedLongLong(o);
  if (v == (unsigned long long)-1 && PyErr_Occurred()) return -1;
  *out = (void*)(uintptr_t)v;
  return 0;
}

/* a C-contiguous uint8 array of exactly (h, w, s) */
static int get_stack(PyObject* o, Py_buffer* view, Py_ssize_t h, Py_ssize_t w, Py_ssize_t s) {
  if (PyObject_GetBuffer(o, view, PyBUF_C_CONTIGUOUS | PyBUF_FORMAT) != 0) { PyErr_Clear(); return 0; }
  int ok = view->itemsize == 1 && view->ndim == 3 && view->shape[0] == h && view->shape[1] == w && view->shape[2] == s &&
           (view->format == NULL || (view->format[0] == 'B' && view->format[1] == 0));
  if (!ok) PyBuffer_Release(view);
  return ok;
}

/* put(fn, handle, slot, s_tm1, s_t, action, reward, discount, stream, H, W, S) -> status */
static PyObject* fc_put(PyObject* self, PyObject* const* a, Py_ssize_t n) {
  (void)self;
  if (n != 12) { PyErr_SetString(PyExc_TypeError, "put takes 12 arguments"); return NULL; }
  void *fn, *h, *stream;
  if (as_ptr(a[0], &fn) || as_ptr(a[1], &h) || as_ptr(a[8], &stream)) return NULL;
  const long long slot = PyLong_AsLongLong(a[2]);
  if (slot == -1 && PyErr_Occurred()) { PyErr_Clear(); return PyLong_FromLong(kFallback); }
  const Py_ssize_t H = PyLong_AsSsize_t(a[9]), W = PyLong_AsSsize_t(a[10]), S = PyLong_AsSsize_t(a[11]);
  if (PyErr_Occurred()) return NULL;
  /* scalars: anything with __index__ / __float__ (Python and NumPy scalars) */
  const long action = PyLong_AsLong(a[5]);
  if (action == -1 && PyErr_Occurred()) { PyErr_Clear(); return PyLong_FromLong(kFallback); }
  const double reward = PyFloat_AsDouble(a[6]);
  if (reward == -1.0 && PyErr_Occurred()) { PyErr_Clear(); return PyLong_FromLong(kFallback); }
  const double discount = PyFloat_AsDouble(a[7]);
  if (discount == -1.0 && PyErr_Occurred()) { PyErr_Clear(); return PyLong_FromLong(kFallback); }
  Py_buffer v0, v1;
  if (!get_stack(a[3], &v0, H, W, S)) return PyLong_FromLong(kFallback);
  if (!get_stack(a[4], &v1, H, W, S)) { PyBuffer_Release(&v0); return PyLong_FromLong(kFallback); }
  const int rc = ((put_fn)fn)(h, (int64_t)slot, (const uint8_t*)v0.buf, (const uint8_t*)v1.buf, (int32_t)action,
                              (float)reward, (float)discount, (uintptr_t)stream);
  PyBuffer_Release(&v0);
  PyBuffer_Release(&v1);
  return PyLong_FromLong(rc);
}

/* update(fn, learner, replay, slots, stream, batch) -> status; slots: C-contiguous int64 of `batch` entries */
static PyObject* fc_update(PyObject* self, PyObject* const* a, Py_ssize_t n) {
  (void)self;
  if (n != 6) { PyErr_SetString(PyExc_TypeError, "update takes 6 arguments"); return NULL; }
  void *fn, *h, *rep, *stream;
  if (as_ptr(a[0], &fn) || as_ptr(a[1], &h) || as_ptr(a[2], &rep) || as_ptr(a[4], &stream)) return NULL;
  const Py_ssize_t batch = PyLong_AsSsize_t(a[5]);
  if (PyErr_Occurred()) return NULL;
  Py_buffer v;
  if (PyObject_GetBuffer(a[3], &v, PyBUF_C_CONTIGUOUS | PyBUF_FORMAT) != 0) { PyErr_Clear(); return PyLong_FromLong(kFallback); }
  const int is_i64 = v.itemsize == 8 && v.format != NULL && (v.format[0] == 'l' || v.format[0] == 'q') && v.format[1] == 0;
  if (!is_i64 || v.len != batch * 8) { PyBuffer_Release(&v); return PyLong_FromLong(kFallback); }
  const int rc = ((update_fn)fn)(h, rep, (const int64_t*)v.buf, (uintptr_t)stream);
  PyBuffer_Release(&v);
  return PyLong_FromLong(rc);
}

/* loss_lagged(post_fn, take_fn, learner, stream, take) -> (status, loss or None): posts a read-back of the latest
 * update's loss; with `take` also waits for the oldest outstanding one */
static PyObject* fc_loss_lagged(PyObject* self, PyObject* const* a, Py_ssize_t n) {
  (void)self;
  if (n != 5) { PyErr_SetString(PyExc_TypeError, "loss_lagged takes 5 arguments"); return NULL; }
  void *post, *take, *h, *stream;
  if (as_ptr(a[0], &post) || as_ptr(a[1], &take) || as_ptr(a[2], &h) || as_ptr(a[3], &stream)) return NULL;
  const int want = PyObject_IsTrue(a[4]);
  if (want < 0) return NULL;
  int rc = ((post_fn)post)(h, (uintptr_t)stream);
  if (rc != 0 || !want) return Py_BuildValue("(iO)", rc, Py_None);
  float loss = 0.f;
  rc = ((take_fn)take)(h, &loss);
  if (rc != 0) return Py_BuildValue("(iO)", rc, Py_None);
  return Py_BuildValue("(id)", 0, (double)loss);
}

static PyMethodDef methods[] = {
    {"put", (PyCFunction)(void (*)(void))fc_put, METH_FASTCALL, "dqnb_replay_put on buffer-protocol stacks"},
    {"update", (PyCFunction)(void (*)(void))fc_update, METH_FASTCALL, "dqnb_learner_update_from_replay on an int64 slot array"},
    {"loss_lagged", (PyCFunction)(void (*)(void))fc_loss_lagged, METH_FASTCALL, "loss_post [+ loss_take]"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef module = {PyModuleDef_HEAD_INIT, "_fastcall", "per-step call helper of dqn_mgsc_zoo_b200", -1, methods,
                                    NULL, NULL, NULL, NULL};

PyMODINIT_FUNC PyInit__fastcall(void) {
  PyObject* m = PyModule_Create(&module);
  if (m != NULL) PyModule_AddIntConstant(m, "FALLBACK", kFallback);
  return m;
}
