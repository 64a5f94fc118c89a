"""Mirror of the deterministic half of reference generate_query/generate_query.py
(:46-57, :88-129, :214-266): metapath -> step text -> Cypher text -> instance rows -> rows whose
destination the event message mentions.  The LLM generator (:18-43, :60-85, :134-211) is out of
scope; its recorded outputs are ordinary Cypher text and go through run_and_filter_query too - the
functions that talk to it are kept as plumbing for the end-to-end driver (names and return shapes)."""
from common.openai_generic_assistant import OpenAIGenericAssistant

FIXED_STEPS = (("HasEvent", "Event", "EVENT", "metadata_uid"),)


def extend_metapath_construct_string(partial_path):
    """Path -> 'relType, srcKind, destKind, key;' lines, preceded by the two event steps (:46-57)."""
    first_kind = partial_path.nodes[0]["kind"]
    steps = list(FIXED_STEPS) + [("ReferInternal", "Event", first_kind, "involvedObject_uid")]
    steps += [(rel.type, rel["srcKind"], rel["destKind"], rel["key"]) for rel in partial_path.relationships]
    return "\n" + "".join("    " + ", ".join(st) + ";\n" for st in steps)


def human_generate_cypher_query(metapath_str, error_message):
    """Deterministic template (:214-266): one alias per KIND in first-seen order (EVENT is `evt`),
    one MATCH per step in its stored direction, aliases and relationships interleaved in RETURN."""
    steps = [part.strip().split(", ") for part in metapath_str.split(";")[:-1]]
    alias = {"EVENT": "evt"}
    for _, src, dst, _ in steps:
        for kind in (src, dst):
            if kind not in alias:
                alias[kind] = "n%d" % len(alias)
    lines = ["MATCH (evt:EVENT)", "WHERE evt.message CONTAINS %r" % error_message, "WITH evt", "LIMIT 1"]
    for i, (rel_type, src, dst, key) in enumerate(steps, start=1):
        lines.append("MATCH (%s:%s)-[r%d:%s]->(%s:%s)" % (alias[src], src, i, rel_type, alias[dst], dst))
        lines.append("WHERE r%d.key = %r" % (i, key))
    nodes = list(alias.values())
    rels = ["r%d" % i for i in range(1, len(steps) + 1)]
    assert len(nodes) == len(rels) + 1
    cols = [None] * (len(nodes) + len(rels))
    cols[0::2], cols[1::2] = nodes, rels
    lines.append("RETURN " + ", ".join(cols))
    query = "\n".join(lines)
    print("the human generated cypher query is: \n%s" % query)
    return query


def run_and_filter_query(query_executor, cypher_query):
    """Rows of the instance query whose destination the event message mentions (:88-102)."""
    kept = [rec for rec in query_executor.run_query(cypher_query) if message_compatible(rec)]
    if not kept:
        print("Warning: ALL records are not message compatible")
    return kept


_NAME_KEY_BY_TAG = {"nfs": "path", "hostPath": "path", "container": "containerName", "image": "imageName"}


def message_compatible(record):
    """The destination (last column) is mentioned by name or by kind in the event message (:104-129).
    Like the reference this raises when the destination has no usable name key."""
    for ele in record:
        if ele["kind"] == "Event":
            message = ele["message"]
    dest = record[len(record) - 1]
    if dest["isNative"] == "true":
        name_key, kind_key = "name2", "kind2"
    else:
        if dest["isAtomic"] == "true":
            name_key = "val"
        elif dest["tag"] in _NAME_KEY_BY_TAG:
            name_key = _NAME_KEY_BY_TAG[dest["tag"]]
        if dest["isNative"] == "false":
            kind_key = "tag"
    return (dest[name_key] in message) or (dest[kind_key] in message)


# ---- LLM plumbing (reference :18-43, :60-85)
def setup_cypher_generator():
    gen = OpenAIGenericAssistant()
    gen.create_assistant("You write Cypher for Neo4j.", "cypher-query-generator", "gpt-4")
    gen.create_thread()
    print(gen.assistant.id)
    print(gen.thread.id)
    gen.add_message("Each line 'relType, srcKind, destKind, key;' of a metapath becomes one MATCH "
                    "(a:srcKind)-[r:relType]->(b:destKind) WHERE r.key = key; the EVENT node is found by its message.")
    return gen


def extract_cypher(message_str):
    return message_str.split("```cypher")[1].split("```")[0].strip()


def generate_cypher_query(metapath_str, error_message, cypherQueryGenerator):
    cypherQueryGenerator.add_message(
        "Write the Cypher query for this metapath, inside a ```cypher block, keeping every (srcKind)-[rel]->(destKind) "
        "in the given direction.\nthe provided metapath is:\n%s\nthe error message to filtering is:\n%s\n"
        % (metapath_str, error_message))
    print("run assistant")
    cypherQueryGenerator.run_assistant()
    query = extract_cypher(cypherQueryGenerator.wait_get_last_k_message(1).data[0].content[0].text.value)
    print("the generated cypher query is :\n %s" % query)
    return query
