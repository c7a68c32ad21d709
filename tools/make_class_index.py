"""Writes distributed-machine-learning_b200/imagenet_class_index.json.

Keras' decode_predictions downloads imagenet_class_index.json on first use
(/root/reference/models.py:42, :67); offline it is unavailable.  Class NAMES in the
standard ILSVRC-2012 index order come from torchvision's metadata (spaces ->
underscores, as Keras prints them); WNIDS are only known for the classes that
appear in the reference's sample outputs (/root/reference/download/output_*.json)
and the Imagenette subset; the rest get the placeholder "n????????".  A real
imagenet_class_index.json can be dropped in through $DML_B200_CLASS_INDEX.
Run once in the build container: python tools/make_class_index.py
"""
import glob
import json
import os

from torchvision.models._meta import _IMAGENET_CATEGORIES as NAMES

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
known = {}
for f in glob.glob("/root/reference/download/output_*.json"):
    for v in json.load(open(f)).values():
        for wnid, name, _ in v[0]:
            known[name] = wnid
imagenette = {"tench": "n01440764", "English_springer": "n02102040", "cassette_player": "n02979186",
              "chain_saw": "n03000684", "church": "n03028079", "French_horn": "n03394916",
              "garbage_truck": "n03417042", "gas_pump": "n03425413", "golf_ball": "n03445777",
              "parachute": "n03888257"}
known.update(imagenette)
out = {}
hits = 0
for i, n in enumerate(NAMES):
    name = n.replace(" ", "_")
    wnid = known.get(name, "n????????")
    hits += wnid != "n????????"
    out[str(i)] = [wnid, name]
path = os.path.join(ROOT, "distributed-machine-learning_b200", "imagenet_class_index.json")
json.dump(out, open(path, "w"), indent=0)
print(path, "known wnids:", hits, "of", len(known))
