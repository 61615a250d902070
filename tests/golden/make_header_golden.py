"""Golden "MGARD" headers, made with the official protobuf runtime.

TEST INFRASTRUCTURE (run in the build container only; /root/reference does not
exist on the GPU box).  The reference's Metadata.cpp needs protoc-generated code
and cannot be compiled here, so the header is pinned one level down:

  * the message schema is read from the reference's own src/mgard.proto (a small
    parser below turns it into a FileDescriptorProto -- field numbers, enum
    values, oneofs and repeated-ness all come from that file);
  * the Header message is filled the way MetadataBase::Serialize does it
    (src/mgard-x/Metadata/Metadata.cpp:268-462 -- restated in fill_header, including
    its quirk of writing the FILE version into mgard_version and leaving
    file_format_version empty) and serialized by google.protobuf (upb), the same
    wire-format implementation family the reference links;
  * preamble = "MGARD" + u64 LE header size + u32 LE zlib.crc32(header)
    (Metadata.cpp:464-483).

Output: tests/golden/header_v1.json  {case name: {fields..., "hex": "..."}}.
"""
import json
import math
import os
import re
import struct
import sys
import zlib

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

HERE = os.path.dirname(os.path.abspath(__file__))
PROTO = "/root/reference/src/mgard.proto"
FD = descriptor_pb2.FieldDescriptorProto
SCALARS = {"uint64": FD.TYPE_UINT64, "double": FD.TYPE_DOUBLE, "bool": FD.TYPE_BOOL,
           "string": FD.TYPE_STRING}


def parse_proto(text):
    text = re.sub(r"//[^\n]*", "", text)
    toks = re.findall(r"[A-Za-z_][A-Za-z0-9_.]*|\d+|[{}=;]", text)
    pos = 0
    fdp = descriptor_pb2.FileDescriptorProto()
    fdp.name = "mgard.proto"
    fdp.syntax = "proto3"

    def take():
        nonlocal pos
        pos += 1
        return toks[pos - 1]

    def parse_enum(parent):
        e = parent.enum_type.add()
        e.name = take()
        assert take() == "{"
        while toks[pos] != "}":
            if toks[pos] == "reserved":
                while take() != ";":
                    pass
                continue
            v = e.value.add()
            v.name = take()
            assert take() == "="
            v.number = int(take())
            assert take() == ";"
        take()

    def parse_field(msg, oneof_index=None):
        label = FD.LABEL_OPTIONAL
        t = take()
        if t == "repeated":
            label = FD.LABEL_REPEATED
            t = take()
        f = msg.field.add()
        f.name = take()
        assert take() == "="
        f.number = int(take())
        assert take() == ";"
        f.label = label
        if t in SCALARS:
            f.type = SCALARS[t]
        else:
            f.type_name = t          # resolved below
        if oneof_index is not None:
            f.oneof_index = oneof_index

    def parse_message(parent, scope):
        m = parent.message_type.add() if parent is fdp else parent.nested_type.add()
        m.name = take()
        full = scope + "." + m.name
        assert take() == "{"
        while toks[pos] != "}":
            if toks[pos] == "enum":
                take()
                parse_enum(m)
            elif toks[pos] == "message":
                take()
                parse_message(m, full)
            elif toks[pos] == "oneof":
                take()
                o = m.oneof_decl.add()
                o.name = take()
                assert take() == "{"
                while toks[pos] != "}":
                    parse_field(m, len(m.oneof_decl) - 1)
                take()
            else:
                parse_field(m)
        take()

    while pos < len(toks):
        t = take()
        if t == "syntax":
            while take() != ";":
                pass
        elif t == "package":
            fdp.package = take()
            take()
        elif t == "message":
            parse_message(fdp, "." + fdp.package)

    # resolve type names: nested enum/message of the same message, else top level
    top = {m.name for m in fdp.message_type}

    def resolve(m, scope):
        full = scope + "." + m.name
        local_enums = {e.name for e in m.enum_type}
        local_msgs = {n.name for n in m.nested_type}
        for f in m.field:
            if f.type_name:
                n = f.type_name
                if n in local_enums:
                    f.type, f.type_name = FD.TYPE_ENUM, full + "." + n
                elif n in local_msgs:
                    f.type, f.type_name = FD.TYPE_MESSAGE, full + "." + n
                else:
                    assert n in top, n
                    f.type, f.type_name = FD.TYPE_MESSAGE, "." + fdp.package + "." + n
        for n in m.nested_type:
            resolve(n, full)

    for m in fdp.message_type:
        resolve(m, "." + fdp.package)
    return fdp


def header_class():
    fdp = parse_proto(open(PROTO).read())
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("mgard.pb.Header"))


FILE_VERSION = (1, 0, 0)      # CMakeLists.txt:17-19


def fill_header(Header, c):
    """MetadataBase::Serialize, Metadata.cpp:268-462, for a compression header."""
    h = Header()
    v = h.mgard_version
    v.major_, v.minor_, v.patch_ = FILE_VERSION          # :287-291 (sic)
    h.file_format_version.SetInParent()                   # :283-284, never filled
    d = h.domain
    d.topology = 0
    g = d.cartesian_grid_topology
    g.dimension = len(c["shape"])
    g.shape.extend(c["shape"])
    d.geometry = 0                                        # UNIT_CUBE
    h.dataset.type = 1 if c["dtype"] == "f64" else 0
    h.dataset.dimension = 1
    e = h.error_control
    if c["mode"] == "ABS":
        e.mode = 0
    else:
        e.mode = 1
        e.norm_of_original_data = c["norm"]
    e.norm = 0 if math.isinf(c["s"]) else 1
    e.s = c["s"]
    e.tolerance = c["tol"]
    dd = h.domain_decomposition
    dd.method = 1 if c["decomposed"] else 0               # MaxDim only
    dd.decomposition_dimension = c["dd_dim"]
    dd.decomposition_size = c["dd_size"]
    fd = h.function_decomposition
    fd.transform = 0
    fd.hierarchy = 1                                      # MULTIDIMENSION_WITH_GHOST_NODES
    fd.L_target = 0                                       # never filled on this path
    q = h.quantization
    q.method, q.bin_widths, q.type, q.big_endian = 1, 0, 3, False
    h.bitplane_encoding.method = 0
    h.bitplane_encoding.SetInParent()
    en = h.encoding
    en.preprocessor = 0
    # X_HUFFMAN, or X_HUFFMAN_ZSTD for lossless_type::Huffman_Zstd (:431-444; both
    # carry the dictionary and block sizes); the numbers come from the .proto
    comp = en.DESCRIPTOR.fields_by_name["compressor"].enum_type
    en.compressor = comp.values_by_name[
        "X_HUFFMAN_ZSTD" if c.get("lossless") == "zstd" else "X_HUFFMAN"].number
    en.huffman_dictionary_size = c["dict"]
    en.huffman_block_size = c["block"]
    h.device.backend = 3                                  # X_CUDA
    return h


def f32(x):
    return struct.unpack("<f", struct.pack("<f", x))[0]


CASES = {
    # FillForCompression casts tol, s and norm to T first (Metadata.hpp:79-112)
    "f32_3d_rel_linf": dict(shape=[33, 17, 65], dtype="f32", mode="REL", s=math.inf,
                            tol=f32(1e-3), norm=f32(0.987654321), decomposed=False,
                            dd_dim=0, dd_size=33, dict=8192, block=1024),
    "f64_2d_abs_l2": dict(shape=[129, 257], dtype="f64", mode="ABS", s=0.0, tol=1e-4,
                          norm=1.0, decomposed=False, dd_dim=0, dd_size=129,
                          dict=8192, block=1024),
    "f32_3d_rel_decomposed": dict(shape=[1025, 1025, 1025], dtype="f32", mode="REL",
                                  s=math.inf, tol=f32(1e-3), norm=f32(1.05),
                                  decomposed=True, dd_dim=0, dd_size=513,
                                  dict=8192, block=1024),
    "f64_5d_rel_snorm": dict(shape=[8, 16, 39, 39, 39969], dtype="f64", mode="REL",
                             s=1.0, tol=1e-5, norm=3.25, decomposed=True, dd_dim=4,
                             dd_size=4997, dict=4096, block=2048),
    "f32_1d_abs_linf": dict(shape=[100000], dtype="f32", mode="ABS", s=math.inf,
                            tol=f32(0.5), norm=1.0, decomposed=False, dd_dim=0,
                            dd_size=100000, dict=64, block=256),
    "f32_3d_rel_zstd": dict(shape=[257, 130, 65], dtype="f32", mode="REL", s=math.inf,
                            tol=f32(1e-4), norm=f32(2.5), decomposed=True, dd_dim=0,
                            dd_size=129, dict=8192, block=1024, lossless="zstd"),
    "f64_4d_abs_negs": dict(shape=[5, 6, 7, 300], dtype="f64", mode="ABS", s=-0.5,
                            tol=2.5e-7, norm=1.0, decomposed=True, dd_dim=3,
                            dd_size=150, dict=8192, block=1024),
}


def main():
    Header = header_class()
    out = {}
    for name, c in CASES.items():
        body = fill_header(Header, c).SerializeToString(deterministic=True)
        blob = (b"MGARD" + struct.pack("<Q", len(body)) +
                struct.pack("<I", zlib.crc32(body) & 0xFFFFFFFF) + body)
        rec = dict(c)
        rec["s"] = "inf" if math.isinf(c["s"]) else c["s"]
        rec["hex"] = blob.hex()
        out[name] = rec
    path = os.path.join(HERE, "header_v1.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote", path, {k: len(v["hex"]) // 2 for k, v in out.items()})


if __name__ == "__main__":
    sys.exit(main())
