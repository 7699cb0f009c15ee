"""Wire types of the `SemanticSegmentation` gRPC service (sem_seg_server.proto:1-49), built at import
time from a descriptor assembled field by field -- no protoc, no generated *_pb2.py (the reference's
generated module predates protobuf 4 and does not import under current protobuf).

The serialized descriptor is checked byte for byte against the reference's in
tests/test_server.py, so messages produced here are wire-identical.
"""
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

PACKAGE = "sem_seg_server"
SERVICE = "SemanticSegmentation"

_F = descriptor_pb2.FieldDescriptorProto


def _field(msg, name, number, ftype, label=_F.LABEL_OPTIONAL, type_name=None):
    f = msg.field.add()
    f.name, f.number, f.type, f.label = name, number, ftype, label
    if type_name:
        f.type_name = type_name
    return f


def build_file_descriptor_proto():
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "sem_seg_server.proto"
    fd.package = PACKAGE
    fd.syntax = "proto3"

    obj = fd.message_type.add()                                   # sem_seg_server.proto:14-36
    obj.name = "SegmentedObject"
    _field(obj, "label", 1, _F.TYPE_STRING)
    _field(obj, "score", 2, _F.TYPE_FLOAT)
    _field(obj, "area", 3, _F.TYPE_FLOAT)
    _field(obj, "centroid", 4, _F.TYPE_MESSAGE, type_name=".%s.SegmentedObject.Centroid" % PACKAGE)
    cen = obj.nested_type.add()                                   # :30-33
    cen.name = "Centroid"
    _field(cen, "cx", 1, _F.TYPE_FLOAT)
    _field(cen, "cy", 2, _F.TYPE_FLOAT)

    data = fd.message_type.add()                                  # :38-41
    data.name = "SegmentedObjectData"
    _field(data, "data", 1, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, ".%s.SegmentedObject" % PACKAGE)

    res = fd.message_type.add()                                   # :43-46
    res.name = "CameraResolution"
    _field(res, "width", 1, _F.TYPE_INT32)
    _field(res, "height", 2, _F.TYPE_INT32)

    fd.message_type.add().name = "Empty"                          # :48-49

    svc = fd.service.add()                                        # :8-12
    svc.name = SERVICE
    for name, out in (("GetSegmentedObjects", "SegmentedObjectData"), ("GetCameraResolution", "CameraResolution")):
        m = svc.method.add()
        m.name = name
        m.input_type = ".%s.Empty" % PACKAGE
        m.output_type = ".%s.%s" % (PACKAGE, out)
        m.options.SetInParent()        # the reference's descriptor carries empty method options ("{}" in the .proto)
    return fd


FILE_DESCRIPTOR_PROTO = build_file_descriptor_proto()
_pool = descriptor_pool.DescriptorPool()
_file = _pool.Add(FILE_DESCRIPTOR_PROTO) if hasattr(_pool, "Add") else None
if _file is None:
    _pool.AddSerializedFile(FILE_DESCRIPTOR_PROTO.SerializeToString())


def _cls(name):
    return message_factory.GetMessageClass(_pool.FindMessageTypeByName("%s.%s" % (PACKAGE, name)))


SegmentedObject = _cls("SegmentedObject")
SegmentedObjectData = _cls("SegmentedObjectData")
CameraResolution = _cls("CameraResolution")
Empty = _cls("Empty")

METHOD_GET_OBJECTS = "/%s.%s/GetSegmentedObjects" % (PACKAGE, SERVICE)
METHOD_GET_RESOLUTION = "/%s.%s/GetCameraResolution" % (PACKAGE, SERVICE)
