// _tkpack -- the one piece of the library that includes Python.h (optional pack
// helper of the C-ABI contract): walks the list of ConsumerRecord
// objects that consumer.poll() returned and harvests, per record, the address and
// length of its `value` bytes plus its offset into caller-provided arrays.  The
// payload copy itself is done afterwards by tk_slot_pack() (libtorchkafka_b200.so)
// on several threads with the GIL released.
//
// This replaces the per-record attribute access of the reference's hot loop
// (`for record in self._consumer: ... record.value`, src/kafka_dataset.py:156-159)
// with ~20 ns of C per record.
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>

// harvest(records, value_idx, offset_idx, first, max_records, max_bytes,
//         ptrs_addr, lens_addr, offs_addr) -> (taken, total_bytes, keepalive_or_None)
//   records     list/tuple of record objects (namedtuples take the index fast path;
//               value_idx/offset_idx < 0 => attribute lookup by name)
//   first       index of the first record to take
//   *_addr      addresses of uint64[max_records], uint32[max_records], int64[max_records]
static PyObject* harvest(PyObject*, PyObject* args) {
    PyObject* records;
    Py_ssize_t vi, oi, first, max_records;
    unsigned long long max_bytes, ptrs_addr, lens_addr, offs_addr;
    if (!PyArg_ParseTuple(args, "OnnnnKKKK", &records, &vi, &oi, &first, &max_records, &max_bytes,
                          &ptrs_addr, &lens_addr, &offs_addr))
        return nullptr;
    PyObject* seq = PySequence_Fast(records, "records must be a sequence");
    if (!seq) return nullptr;
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(seq);
    PyObject** items = PySequence_Fast_ITEMS(seq);
    uint64_t* ptrs = reinterpret_cast<uint64_t*>(ptrs_addr);
    uint32_t* lens = reinterpret_cast<uint32_t*>(lens_addr);
    int64_t* offs = reinterpret_cast<int64_t*>(offs_addr);
    PyObject* keep = nullptr;      // owns temporaries of the slow paths
    static PyObject* s_value = nullptr;
    static PyObject* s_offset = nullptr;
    if (!s_value) { s_value = PyUnicode_InternFromString("value"); s_offset = PyUnicode_InternFromString("offset"); }

    Py_ssize_t taken = 0;
    unsigned long long total = 0;
    const bool fast = vi >= 0;
    for (Py_ssize_t i = first; i < n && taken < max_records; ++i) {
        PyObject* rec = items[i];
        // Each record is three dependent cache misses (record tuple -> value bytes header,
        // offset int).  Prefetch the tuple 16 records ahead and its two fields 8 ahead.
        if (fast) {
            if (i + 16 < n) __builtin_prefetch(items[i + 16]);
            if (i + 8 < n) {
                PyObject* r8 = items[i + 8];
                if (PyTuple_Check(r8) && PyTuple_GET_SIZE(r8) > vi && PyTuple_GET_SIZE(r8) > oi) {
                    __builtin_prefetch(PyTuple_GET_ITEM(r8, vi));
                    __builtin_prefetch(PyTuple_GET_ITEM(r8, oi));
                }
            }
        }
        PyObject *val, *off;
        bool own = false;
        if (vi >= 0 && PyTuple_Check(rec) && PyTuple_GET_SIZE(rec) > vi && PyTuple_GET_SIZE(rec) > oi) {
            val = PyTuple_GET_ITEM(rec, vi);
            off = PyTuple_GET_ITEM(rec, oi);
        } else {
            val = PyObject_GetAttr(rec, s_value);
            if (!val) goto fail;
            off = PyObject_GetAttr(rec, s_offset);
            if (!off) { Py_DECREF(val); goto fail; }
            own = true;
        }
        if (val == Py_None) {                      // tombstone: empty payload
            if (own) { Py_DECREF(val); }
            val = nullptr;
        } else if (!PyBytes_Check(val)) {          // bytearray / memoryview / ...: materialise once
            PyObject* b = PyBytes_FromObject(val);
            if (own) Py_DECREF(val);
            if (!b) { if (own) Py_DECREF(off); goto fail; }
            if (!keep && !(keep = PyList_New(0))) { Py_DECREF(b); if (own) Py_DECREF(off); goto fail; }
            PyList_Append(keep, b);
            Py_DECREF(b);
            val = b;
        } else if (own) {
            if (!keep && !(keep = PyList_New(0))) { Py_DECREF(val); Py_DECREF(off); goto fail; }
            PyList_Append(keep, val);
            Py_DECREF(val);
        }
        {
            const Py_ssize_t len = val ? PyBytes_GET_SIZE(val) : 0;
            if (total + (unsigned long long)len > max_bytes) { if (own) Py_DECREF(off); break; }
            const long long o = PyLong_AsLongLong(off);
            if (own) Py_DECREF(off);
            if (o == -1 && PyErr_Occurred()) goto fail;
            ptrs[taken] = val ? (uint64_t)(uintptr_t)PyBytes_AS_STRING(val) : 0;
            lens[taken] = (uint32_t)len;
            offs[taken] = o;
            total += (unsigned long long)len;
            ++taken;
        }
    }
    Py_DECREF(seq);
    {
        PyObject* r = Py_BuildValue("(nKO)", taken, total, keep ? keep : Py_None);
        Py_XDECREF(keep);
        return r;
    }
fail:
    Py_DECREF(seq);
    Py_XDECREF(keep);
    return nullptr;
}

static PyMethodDef methods[] = {
    {"harvest", harvest, METH_VARARGS, "Collect (address, length, offset) of every record's value."},
    {nullptr, nullptr, 0, nullptr}};

static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_tkpack", "torchkafka_b200 pack helper", -1, methods};

PyMODINIT_FUNC PyInit__tkpack(void) { return PyModule_Create(&moddef); }
