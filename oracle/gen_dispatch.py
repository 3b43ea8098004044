"""sbt.json -> callHit / callMiss / callAnyHit switch source (row index = case label), for building
user programs of radiance-ray-tracing_b200/programs/ against the REFERENCE runtime in oracle/_ref.
Same table semantics as tools/genSBT.py:53-82 with the current call shapes (radiance.cl:23-28)."""
import json
import sys

rows = json.load(open(sys.argv[1]))
tail = "struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler"
out = ["void callHit(int sbtRecordOffset, struct Payload* payload, struct HitData* hitData, %s)\n{" % tail,
       "    switch ((int)(hitData->instanceSBTOffset + sbtRecordOffset)) {"]
out += ["        case %d: %s(payload, hitData, sceneData, imageArray, sampler); break;" % (i, r["closestHit"])
        for i, r in enumerate(rows) if r.get("closestHit")]
out += ["    }\n}", "void callMiss(int missIndex, struct Payload* payload, %s)\n{" % tail, "    switch (missIndex) {"]
out += ["        case %d: %s(payload, sceneData, imageArray, sampler); break;" % (i, r["miss"])
        for i, r in enumerate(rows) if r.get("miss")]
out += ["    }\n}", "void callAnyHit(bool* cont, int sbtRecordOffset, struct Payload* payload, struct HitData* hitData, %s)\n{" % tail,
        "    switch ((int)(hitData->instanceSBTOffset + sbtRecordOffset)) {"]
out += ["        case %d: %s(cont, payload, hitData, sceneData, imageArray, sampler); break;" % (i, r["anyHit"])
        for i, r in enumerate(rows) if r.get("anyHit")]
out += ["    }\n}"]
print("\n".join(out))
