"""The API reference under doc/ is generated from the sources (scripts/gen_api_doc.py; the reference builds its
pages with doxygen + sphinx, doc/Doxyfile, doc/conf.py): it must be current and must cover the exported surface."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_generated_api_reference_is_current():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "gen_api_doc.py"), "--check"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    assert r.returncode == 0, r.stdout.decode()


def test_reference_pages_cover_the_public_surface(built):
    c_doc = open(os.path.join(ROOT, "doc", "c_api.md")).read()
    # every Rabit* symbol of the reference's C ABI and every device entry point of the header
    for name in ["RabitInit", "RabitFinalize", "RabitGetRingPrevRank", "RabitGetRank", "RabitGetWorldSize", "RabitIsDistributed",
                 "RabitTrackerPrint", "RabitGetProcessorName", "RabitBroadcast", "RabitAllgather", "RabitAllreduce",
                 "RabitLoadCheckPoint", "RabitCheckPoint", "RabitVersionNumber", "RabitLinkTag"]:
        assert re.search(r"\b%s\(" % name, c_doc), name
    header = open(os.path.join(ROOT, "include", "rabit_device.h")).read()
    for name in set(re.findall(r"\b(Rabit(?:Device|Emu)\w+)\(", header)):
        assert name + "(" in c_doc, name
    cpp_doc = open(os.path.join(ROOT, "doc", "cpp_api.md")).read()
    for name in ["Init(", "Finalize(", "Allreduce(", "Broadcast(", "Allgather(", "LoadCheckPoint(", "CheckPoint(", "LazyCheckPoint(",
                 "VersionNumber(", "class Reducer", "class SerializeReducer", "class IEngine", "class Serializable", "class Stream"]:
        assert name in cpp_doc, name
    py_doc = open(os.path.join(ROOT, "doc", "python_api.md")).read()
    for name in ["init(", "finalize(", "get_rank(", "get_world_size(", "tracker_print(", "get_processor_name(", "broadcast(",
                 "allreduce(", "load_checkpoint(", "checkpoint(", "version_number(", "allreduce_async_(", "launch("]:
        assert "`" + name in py_doc, name
