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
#include <vector>

struct Sink {
    uint64_t* ptrs;
    uint32_t* lens;
    int64_t* offs;
    Py_ssize_t taken;            // records written so far
    Py_ssize_t max_records;
    unsigned long long total;    // payload bytes so far
    unsigned long long max_bytes;
    PyObject* keep;              // owns temporaries of the slow paths (created on demand)
};

static PyObject* s_value = nullptr;
static PyObject* s_offset = nullptr;

// Append records of `seq` (from index `first`) to the sink until it is full.
// Returns the number of records taken from this sequence, or -1 with an exception set.
static Py_ssize_t harvest_into(Sink& k, PyObject* seq, Py_ssize_t vi, Py_ssize_t oi, Py_ssize_t first) {
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(seq);
    PyObject** items = PySequence_Fast_ITEMS(seq);
    const bool fast = vi >= 0;
    Py_ssize_t i = first;
    for (; i < n && k.taken < k.max_records; ++i) {
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
        if (fast && PyTuple_Check(rec) && PyTuple_GET_SIZE(rec) > vi && PyTuple_GET_SIZE(rec) > oi) {
            val = PyTuple_GET_ITEM(rec, vi);
            off = PyTuple_GET_ITEM(rec, oi);
        } else {
            val = PyObject_GetAttr(rec, s_value);
            if (!val) return -1;
            off = PyObject_GetAttr(rec, s_offset);
            if (!off) { Py_DECREF(val); return -1; }
            own = true;
        }
        if (val == Py_None) {                      // tombstone: empty payload
            if (own) Py_DECREF(val);
            val = nullptr;
        } else if (!PyBytes_Check(val)) {          // bytearray / memoryview / ...: materialise once
            PyObject* b = PyBytes_FromObject(val);
            if (own) Py_DECREF(val);
            if (!b) { if (own) Py_DECREF(off); return -1; }
            if (!k.keep && !(k.keep = PyList_New(0))) { Py_DECREF(b); if (own) Py_DECREF(off); return -1; }
            PyList_Append(k.keep, b);
            Py_DECREF(b);
            val = b;
        } else if (own) {
            if (!k.keep && !(k.keep = PyList_New(0))) { Py_DECREF(val); Py_DECREF(off); return -1; }
            PyList_Append(k.keep, val);
            Py_DECREF(val);
        }
        const Py_ssize_t len = val ? PyBytes_GET_SIZE(val) : 0;
        if (k.total + (unsigned long long)len > k.max_bytes) { if (own) Py_DECREF(off); break; }
        const long long o = PyLong_AsLongLong(off);
        if (own) Py_DECREF(off);
        if (o == -1 && PyErr_Occurred()) return -1;
        k.ptrs[k.taken] = val ? (uint64_t)(uintptr_t)PyBytes_AS_STRING(val) : 0;
        k.lens[k.taken] = (uint32_t)len;
        k.offs[k.taken] = o;
        k.total += (unsigned long long)len;
        ++k.taken;
    }
    return i - first;
}

static void intern_names() {
    if (!s_value) { s_value = PyUnicode_InternFromString("value"); s_offset = PyUnicode_InternFromString("offset"); }
}

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
    intern_names();
    PyObject* seq = PySequence_Fast(records, "records must be a sequence");
    if (!seq) return nullptr;
    Sink k = {reinterpret_cast<uint64_t*>(ptrs_addr), reinterpret_cast<uint32_t*>(lens_addr),
              reinterpret_cast<int64_t*>(offs_addr), 0, max_records, 0, max_bytes, nullptr};
    const Py_ssize_t got = harvest_into(k, seq, vi, oi, first);
    Py_DECREF(seq);
    if (got < 0) { Py_XDECREF(k.keep); return nullptr; }
    PyObject* r = Py_BuildValue("(nKO)", k.taken, k.total, k.keep ? k.keep : Py_None);
    Py_XDECREF(k.keep);
    return r;
}

// ---------------------------------------------------------------- GIL-free fast path
// The records of a poll() result are immutable objects (namedtuples of bytes / int) kept alive by
// the lists we hold, so their fields can be read without the GIL -- which is what lets the thread
// that consumes batches run Python while a producer thread chases 3 pointers per record.  Only
// plain memory reads happen in here: no reference counting, no allocation, no error state.  The
// first record that is not an exact (tuple-subclass, bytes value, int offset) ends the fast path;
// the caller finishes the rest under the GIL.
#if PY_VERSION_HEX >= 0x030C0000
static inline bool long_to_i64_nogil(PyObject* o, int64_t* out) {
    const PyLongObject* l = reinterpret_cast<const PyLongObject*>(o);
    if (_PyLong_IsCompact(l)) { *out = (int64_t)_PyLong_CompactValue(l); return true; }
    const uintptr_t tag = l->long_value.lv_tag;
    const Py_ssize_t nd = (Py_ssize_t)(tag >> 3);
    const int sign = (int)(tag & 3);                 // 0 positive, 1 zero, 2 negative
    if (nd > 3) return false;
    unsigned __int128 v = 0;
    for (Py_ssize_t i = nd; i-- > 0;) v = (v << PyLong_SHIFT) | l->long_value.ob_digit[i];
    if (v > (unsigned __int128)INT64_MAX) return false;
    *out = sign == 2 ? -(int64_t)v : (int64_t)v;
    return true;
}

// returns the number of records consumed from items[first..n); *stopped is set when the sink is
// full (records / bytes) and *bail when a record needs the general path
#ifndef PF_FAR
#define PF_FAR 32
#define PF_NEAR 16
#endif
static Py_ssize_t harvest_nogil(Sink& k, PyObject** items, Py_ssize_t n, Py_ssize_t vi, Py_ssize_t oi,
                                bool* bail) {
    Py_ssize_t i = 0;
    for (; i < n && k.taken < k.max_records; ++i) {
        if (i + PF_FAR < n) {
            const char* far = reinterpret_cast<const char*>(items[i + PF_FAR]);
            __builtin_prefetch(far);
            __builtin_prefetch(far + 64);          // ob_item[vi] of a 12-field record lies in the tuple's second line
        }
        if (i + PF_NEAR < n) {
            PyObject* r8 = items[i + PF_NEAR];
            if (PyType_HasFeature(Py_TYPE(r8), Py_TPFLAGS_TUPLE_SUBCLASS) && Py_SIZE(r8) > vi && Py_SIZE(r8) > oi) {
                __builtin_prefetch(reinterpret_cast<PyTupleObject*>(r8)->ob_item[vi]);
                __builtin_prefetch(reinterpret_cast<PyTupleObject*>(r8)->ob_item[oi]);
            }
        }
        PyObject* rec = items[i];
        if (!PyType_HasFeature(Py_TYPE(rec), Py_TPFLAGS_TUPLE_SUBCLASS) || Py_SIZE(rec) <= vi || Py_SIZE(rec) <= oi) {
            *bail = true; break;
        }
        PyObject* val = reinterpret_cast<PyTupleObject*>(rec)->ob_item[vi];
        PyObject* off = reinterpret_cast<PyTupleObject*>(rec)->ob_item[oi];
        if (!PyType_HasFeature(Py_TYPE(val), Py_TPFLAGS_BYTES_SUBCLASS) ||
            !PyType_HasFeature(Py_TYPE(off), Py_TPFLAGS_LONG_SUBCLASS)) { *bail = true; break; }
        int64_t o;
        if (!long_to_i64_nogil(off, &o)) { *bail = true; break; }
        const Py_ssize_t len = Py_SIZE(val);
        if (k.total + (unsigned long long)len > k.max_bytes) break;
        k.ptrs[k.taken] = (uint64_t)(uintptr_t)reinterpret_cast<PyBytesObject*>(val)->ob_sval;
        k.lens[k.taken] = (uint32_t)len;
        k.offs[k.taken] = o;
        k.total += (unsigned long long)len;
        ++k.taken;
    }
    return i;
}
#define TK_HAVE_NOGIL_HARVEST 1
#else
#define TK_HAVE_NOGIL_HARVEST 0
#endif

// harvest_groups(groups, dense, value_idx, offset_idx, max_records, max_bytes,
//                ptrs_addr, lens_addr, offs_addr, parts_addr)
//     -> (taken, total_bytes, groups_done, taken_in_last_group, keepalive_or_None)
//   groups   list of (topic_partition, records) pairs -- list(poll_result.items())
//   dense    list of int: dense partition index of each group (written to parts[] per record)
// Processes whole poll results in one call.  Stops when the sink is full: `groups_done` groups
// were taken completely and `taken_in_last_group` records of group number `groups_done`.
static PyObject* harvest_groups(PyObject*, PyObject* args) {
    PyObject *groups, *dense;
    Py_ssize_t vi, oi, max_records;
    unsigned long long max_bytes, ptrs_addr, lens_addr, offs_addr, parts_addr;
    if (!PyArg_ParseTuple(args, "OOnnnKKKKK", &groups, &dense, &vi, &oi, &max_records, &max_bytes,
                          &ptrs_addr, &lens_addr, &offs_addr, &parts_addr))
        return nullptr;
    intern_names();
    if (!PyList_Check(groups) || !PyList_Check(dense) || PyList_GET_SIZE(groups) != PyList_GET_SIZE(dense)) {
        PyErr_SetString(PyExc_TypeError, "groups and dense must be lists of equal length");
        return nullptr;
    }
    int32_t* parts = reinterpret_cast<int32_t*>(parts_addr);
    Sink k = {reinterpret_cast<uint64_t*>(ptrs_addr), reinterpret_cast<uint32_t*>(lens_addr),
              reinterpret_cast<int64_t*>(offs_addr), 0, max_records, 0, max_bytes, nullptr};
    const Py_ssize_t ng = PyList_GET_SIZE(groups);
    Py_ssize_t g = 0, partial = 0;
    // under the GIL: pin every group's record sequence and read its dense partition index
    struct Group { PyObject* seq; PyObject** items; Py_ssize_t size; long dense; };
    std::vector<Group> gs;
    gs.reserve((size_t)ng);
    auto release_groups = [&]() { for (auto& x : gs) Py_DECREF(x.seq); };
    for (Py_ssize_t q = 0; q < ng; ++q) {
        PyObject* pair = PyList_GET_ITEM(groups, q);
        if (!PyTuple_Check(pair) || PyTuple_GET_SIZE(pair) != 2) {
            PyErr_SetString(PyExc_TypeError, "each group must be a (partition, records) pair");
            release_groups();
            return nullptr;
        }
        const long d = PyLong_AsLong(PyList_GET_ITEM(dense, q));
        if (d == -1 && PyErr_Occurred()) { release_groups(); return nullptr; }
        PyObject* seq = PySequence_Fast(PyTuple_GET_ITEM(pair, 1), "records must be a sequence");
        if (!seq) { release_groups(); return nullptr; }
        gs.push_back({seq, PySequence_Fast_ITEMS(seq), PySequence_Fast_GET_SIZE(seq), d});
    }
    bool done = false;
#if TK_HAVE_NOGIL_HARVEST
    if (vi >= 0 && oi >= 0) {
        bool bail = false;
        Py_ssize_t bail_at = 0;
        Py_BEGIN_ALLOW_THREADS
        for (; g < ng; ++g) {
            const Py_ssize_t before = k.taken;
            const Py_ssize_t got = harvest_nogil(k, gs[g].items, gs[g].size, vi, oi, &bail);
            for (Py_ssize_t j = before; j < k.taken; ++j) parts[j] = (int32_t)gs[g].dense;
            if (got < gs[g].size) { partial = got; bail_at = got; break; }
        }
        Py_END_ALLOW_THREADS
        if (!bail) done = true;              // all groups taken, or the sink is full
        else {
            // finish group g from record bail_at under the GIL (general path), then go on below
            const Py_ssize_t before = k.taken;
            const Py_ssize_t got = harvest_into(k, gs[g].seq, vi, oi, bail_at);
            if (got < 0) { release_groups(); Py_XDECREF(k.keep); return nullptr; }
            for (Py_ssize_t j = before; j < k.taken; ++j) parts[j] = (int32_t)gs[g].dense;
            if (bail_at + got < gs[g].size) { partial = bail_at + got; done = true; }
            else { partial = 0; ++g; }
        }
    }
#endif
    if (!done) {
        for (; g < ng; ++g) {
            const Py_ssize_t before = k.taken;
            const Py_ssize_t got = harvest_into(k, gs[g].seq, vi, oi, 0);
            if (got < 0) { release_groups(); Py_XDECREF(k.keep); return nullptr; }
            for (Py_ssize_t j = before; j < k.taken; ++j) parts[j] = (int32_t)gs[g].dense;
            if (got < gs[g].size) { partial = got; break; }
        }
    }
    release_groups();
    PyObject* r = Py_BuildValue("(nKnnO)", k.taken, k.total, g, partial, k.keep ? k.keep : Py_None);
    Py_XDECREF(k.keep);
    return r;
}

static PyMethodDef methods[] = {
    {"harvest", harvest, METH_VARARGS, "Collect (address, length, offset) of every record's value."},
    {"harvest_groups", harvest_groups, METH_VARARGS, "harvest() over a whole poll() result."},
    {nullptr, nullptr, 0, nullptr}};

static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_tkpack", "torchkafka_b200 pack helper", -1, methods};

PyMODINIT_FUNC PyInit__tkpack(void) { return PyModule_Create(&moddef); }
