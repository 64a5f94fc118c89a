"""Mirror of the graph half of reference find_metapath/find_srckind_metapath_neo4j.py:63-175.

Same function names, arguments, return values and failure behaviour; the statements sent to the
executor have the semantics of the reference's (:64-68, :76-84, :95-129).  The LLM half of that module (:20-60, :178-240) is out of
scope: what is here of it is the plumbing the end-to-end driver needs - same
function names and return shapes, prompts reduced to what a recorded-response assistant keys on."""
import json

from common.openai_generic_assistant import OpenAIGenericAssistant

KINDS_QUERY = (
    "MATCH (n1) WHERE n1.category IN ['NativeEntity', 'ExternalEntity'] "
    "RETURN n1.category AS category, n1.kind AS kind")

SRC_KIND_QUERY = (
    "MATCH (n1:Event)-[s1:HasEvent]->(N1:EVENT) WHERE N1.message CONTAINS $message "
    "WITH n1, N1, s1 "
    "MATCH (n1:Event)-[r1:ReferInternal]->(n2) WHERE r1.key = 'involvedObject_uid' "
    "RETURN DISTINCT n2.kind2 LIMIT 5")

_PATH_FILTER = (
    "WHERE n1.kind = $srcKind AND n2.kind = $destKind "
    "AND all(node IN nodes(path) WHERE single(x IN nodes(path) WHERE x = node)) "
    "AND all(node IN nodes(path) WHERE NOT node.kind IN ['Event', 'Namespace']) "
    "AND ($intermediateKinds IS NULL OR size($intermediateKinds) = 0 "
    "OR any(node IN nodes(path)[1..-1] WHERE node.kind IN $intermediateKinds)) "
    "RETURN path")

# the reference's preference order (:139-150): directed, undirected, one undirected hop, via Namespace
METAPATH_LADDER = (
    ("directed", "MATCH path = (n1)-[*1..3]->(n2) " + _PATH_FILTER),
    ("undirected", "MATCH path = (n1)-[*1..3]-(n2) " + _PATH_FILTER),
    ("one hop", "MATCH path = (n1)-[r1]-(n2) WHERE n1.kind = $srcKind AND n2.kind = $destKind RETURN path"),
    ("via Namespace", "MATCH path = (n1)-[r1]-(n2)-[r2]-(n3) "
                      "WHERE n1.kind = $srcKind AND n2.kind = 'Namespace' AND n3.kind = $destKind RETURN path"),
)


def find_native_external_kinds(query_executor):
    """-> (sorted native kinds, sorted external kinds)   [reference :63-72]"""
    by_cat = {"NativeEntity": [], "ExternalEntity": []}
    for rec in query_executor.run_query(KINDS_QUERY):
        if rec["category"] in by_cat:
            by_cat[rec["category"]].append(rec["kind"])
    return sorted(by_cat["NativeEntity"]), sorted(by_cat["ExternalEntity"])


def find_srcKind(query_executor, message):
    """Kind of the object the event with this message is about; IndexError when no event matches,
    like the reference's records[0] (:75-90)."""
    records = query_executor.run_query(SRC_KIND_QUERY, {"message": message})
    src_kind = records[0]["n2.kind2"]
    print("srcKind = %s" % src_kind)
    return src_kind


def find_metapath(query_executor, srcKind, destKind, intermediateKinds=None):
    """Shortest metapaths srcKind -> destKind, first rung of the ladder that yields any row
    (:93-160).  TypeError when intermediateKinds is None and ValueError when every rung is empty,
    both as in the reference (its default argument is unusable, its min() sees an empty list)."""
    inter = [k for k in intermediateKinds if k != "Namespace"]
    params = {"srcKind": srcKind, "destKind": destKind, "intermediateKinds": inter}
    records = []
    for name, query in METAPATH_LADDER:
        print("Looking for a %s metapath ...\n" % name)
        records = query_executor.run_query(query, params)
        if len(records) > 0:
            break
    shortest = min(len(rec["path"]) for rec in records)
    metapaths = [rec["path"] for rec in records if len(rec["path"]) == shortest]
    for mp in metapaths:
        print_metapath(path=mp)
    return metapaths


def print_metapath(path):
    print("Nodes:")
    for node in path.nodes:
        print(node["kind"])
    print("Relationships:")
    for rel in path.relationships:
        print(rel.type, rel["srcKind"], rel["destKind"], rel["key"])
    print("----------------------------------")


# ---- LLM plumbing (reference :20-60, :178-240): an assistant that names the destination kind
def setup_root_cause_locator():
    locator = OpenAIGenericAssistant()
    locator.create_assistant("You analyse Kubernetes error messages and name the resource kinds involved.",
                             "k8s-root-cause-locator", "gpt-4")
    locator.create_thread()
    print(locator.assistant.id)
    print(locator.thread.id)
    return locator


def build_prompt_template(nativeKinds, externalKinds):
    """Template with the placeholders {involved_object} and {error_message} (reference :200-240)."""
    return ("Known k8s API resource kinds: " + ", ".join(nativeKinds) + "\nKnown external resource kinds: "
            + ", ".join(externalKinds) + "\nThe error below was reported on a {involved_object}. Name the kind that "
            "resolves it (DestinationKind) and the kinds on the way (RelevantResources), both from the lists above, "
            "as JSON inside a ```json block.\n\n{error_message}\n")


def extract_json(message_str):
    return json.loads(message_str.split("```json")[1].split("```")[0].strip())


def find_destKind_relevantResources(errorMessage, srcKind, promptTemplate, rootCauseLocator):
    rootCauseLocator.add_message(promptTemplate.format(error_message=errorMessage, involved_object=srcKind))
    rootCauseLocator.run_assistant()
    return extract_json(rootCauseLocator.wait_get_last_k_message(1).data[0].content[0].text.value)
