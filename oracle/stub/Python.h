/*
 * TEST INFRASTRUCTURE ONLY -- not part of the product.
 *
 * Minimal stand-in for the CPython-2 <Python.h> so that the reference's
 * csrc/workhorse.c compiles UNMODIFIED into oracle/_ref/libkvarq_ref.so
 * (see oracle/Makefile, oracle/ref_harness.c).  Only the symbols that
 * translation unit touches exist here; the module-glue functions
 * (engine_findseqs, initengine, ...) merely have to compile, they are never
 * called.  The two calls on the scan path that matter are
 * PyString_FromStringAndSize + PyList_Append (workhorse.c:437-438), which the
 * harness uses to capture the `hitseqs` bytes.
 */
#ifndef KVQ_STUB_PYTHON_H
#define KVQ_STUB_PYTHON_H

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdarg.h>

typedef struct _stub_object {
    int kind;           /* 1 = string */
    long size;
    char *data;
} PyObject;

typedef PyObject *(*PyCFunction)(PyObject *, PyObject *);
typedef struct { const char *ml_name; PyCFunction ml_meth; int ml_flags; const char *ml_doc; } PyMethodDef;

#define METH_VARARGS  1
#define METH_KEYWORDS 2
#define PyMODINIT_FUNC void
#define Py_BEGIN_ALLOW_THREADS {
#define Py_END_ALLOW_THREADS }
#define Py_RETURN_NONE return (PyObject *) 0

extern PyObject stub_exc_io, stub_exc_runtime, stub_exc_memory, stub_exc_type;
#define PyExc_IOError      (&stub_exc_io)
#define PyExc_RuntimeError (&stub_exc_runtime)
#define PyExc_MemoryError  (&stub_exc_memory)
#define PyExc_TypeError    (&stub_exc_type)

/* implemented in ref_harness.c; renamed so they cannot be interposed by a real libpython
   when the .so is loaded into a Python process (the Makefile also links -Bsymbolic) */
#define PyString_FromStringAndSize stub_PyString_FromStringAndSize
#define PyList_Append stub_PyList_Append
PyObject *PyString_FromStringAndSize(const char *s, long n);
int PyList_Append(PyObject *list, PyObject *item);
void stub_decref(PyObject *o);
#define Py_DECREF(o)  stub_decref((PyObject *)(o))
#define Py_XDECREF(o) stub_decref((PyObject *)(o))

static inline PyObject *PyString_FromString(const char *s) { (void) s; return 0; }
static inline char *PyString_AsString(PyObject *o) { (void) o; return 0; }
static inline long PyString_Size(PyObject *o) { (void) o; return 0; }
static inline int PyString_Check(PyObject *o) { (void) o; return 0; }
static inline PyObject *PyList_New(long n) { (void) n; return 0; }
static inline PyObject *PyTuple_New(long n) { (void) n; return 0; }
static inline int PyTuple_SetItem(PyObject *t, long i, PyObject *o) { (void) t; (void) i; (void) o; return 0; }
static inline PyObject *PyInt_FromLong(long v) { (void) v; return 0; }
static inline long PyInt_AS_LONG(PyObject *o) { (void) o; return 0; }
static inline PyObject *Py_BuildValue(const char *fmt, ...) { (void) fmt; return 0; }
static inline PyObject *PyObject_CallObject(PyObject *f, PyObject *a) { (void) f; (void) a; return 0; }
static inline PyObject *PyObject_GetAttrString(PyObject *o, const char *n) { (void) o; (void) n; return 0; }
static inline int PyObject_SetAttrString(PyObject *o, const char *n, PyObject *v) { (void) o; (void) n; (void) v; return 0; }
static inline int PyArg_ParseTuple(PyObject *a, const char *fmt, ...) { (void) a; (void) fmt; return 0; }
static inline int PyArg_ParseTupleAndKeywords(PyObject *a, PyObject *k, const char *fmt, char **kwl, ...) { (void) a; (void) k; (void) fmt; (void) kwl; return 0; }
static inline int PySequence_Check(PyObject *o) { (void) o; return 0; }
static inline long PySequence_Size(PyObject *o) { (void) o; return 0; }
static inline PyObject *PySequence_GetItem(PyObject *o, long i) { (void) o; (void) i; return 0; }
static inline PyObject *PyErr_NoMemory(void) { return 0; }
static inline void PyErr_SetString(PyObject *e, const char *s) { (void) e; (void) s; }
static inline PyObject *Py_InitModule(const char *n, PyMethodDef *m) { (void) n; (void) m; return 0; }
static inline PyObject *PyImport_ImportModule(const char *n) { (void) n; return 0; }

#endif
