"""End-to-end incident driver: the mirror's counterpart of the reference's test_all.py:18-157.

Per incident the order of calls is the reference's: source kind of the event (S2) -> destination kind and
relevant kinds from the root-cause assistant, asked up to three times with the parse error fed back
(reference :63-83) -> metapaths on the metagraph (ladder) -> per metapath a generated Cypher query, asked up
to three times with the executor's error fed back (:99-122), and the deterministic template when the
attempts are used up or the query returns nothing (:127-131) -> the state check of every row (:133-134).
The progress lines printed are the reference's, so a transcript of this driver and of the reference's can
be laid side by side.

Differences on purpose: the incident messages are an argument (the reference hard-codes ten messages of its
authors' cluster and runs the second one), and the two attempt loops are one function.
"""
import json
import time

import neo4j

from check_state.analyze_root_cause import check_statepath, setup_state_semantic_analyzer
from common.neo4j_query_executor import Neo4jQueryExecutor
from find_metapath.find_srckind_metapath_neo4j import (build_prompt_template, find_destKind_relevantResources,
                                                       find_metapath, find_native_external_kinds, find_srcKind,
                                                       setup_root_cause_locator)
from generate_query.generate_query import (extend_metapath_construct_string, generate_cypher_query,
                                           human_generate_cypher_query, run_and_filter_query, setup_cypher_generator)

METAGRAPH_URI = "bolt://10.1.0.176:7687"          # the addresses the reference's driver opens (test_all.py:21-22);
STATEGRAPH_URI = "bolt://10.1.0.174:7687"         # here they are aliases of snapshots (executor.register_uri_alias)
MAX_ATTEMPTS = 3


def attempt(fn, assistant, what, special, special_name, hint):
    """fn() up to MAX_ATTEMPTS times; an exception is printed and sent to the assistant as the next message.
    Returns (value or None, index of the last attempt made)."""
    for k in range(MAX_ATTEMPTS):
        try:
            return fn(k), k
        except special as e:
            print(f"{special_name} occurred: {str(e)}")
            assistant.add_message(f"The {what} encounters the following exception:\n{special_name} occurred: {str(e)}\n{hint}")
        except Exception as e:
            print(f"An unexpected error occurred: {str(e)}")
            assistant.add_message(f"The {what} encounters the following exception:\nAn unexpected error occurred: {str(e)}\n"
                                  f"Based on the exception details above, please generate a correct {what}.")
    return None, MAX_ATTEMPTS - 1


def main(messages=(), metagraph_uri=METAGRAPH_URI, stategraph_uri=STATEGRAPH_URI):
    print("create executor and init connection")
    meta = Neo4jQueryExecutor(metagraph_uri, "neo4j", "")
    state = Neo4jQueryExecutor(stategraph_uri, "neo4j", "")
    print("create openai client with assistant and thread")
    print("setup root_cause_locator")
    locator = setup_root_cause_locator()
    print("find native and external kinds and build prompt template")
    native, external = find_native_external_kinds(meta)
    template = build_prompt_template(native, external)
    print("setup cypher_generator")
    generator = setup_cypher_generator()
    print("setup state_semantic_analyzer")
    analyzer = setup_state_semantic_analyzer()
    started = time.time()
    results = []
    for message in messages:
        print(message)
        print("test find_srcKind()")
        src = find_srcKind(state, message)

        def ask_destination(k):
            print("$" * 100)
            print("test find destKind and relevantResources")
            return find_destKind_relevantResources(message, src, template, locator)
        answer, _ = attempt(ask_destination, locator, "dest_relevant", json.decoder.JSONDecodeError, "JSON Error",
                            "make sure to return the output in JSON format, and put it in ```json <dest_relevant> ```")
        print("test find_metapath()")
        dest = answer["DestinationKind"]
        inter = [x for x in answer["RelevantResources"] if x not in (src, dest) and (x in native or x in external)]
        for metapath in find_metapath(meta, src, dest, inter):
            steps = extend_metapath_construct_string(metapath)

            def ask_query(k):
                print("%" * 100)
                print(f"attempt = {k}\n")
                print(f"generate cypher query for the following extended metapath: \n {steps}")
                return run_and_filter_query(state, generate_cypher_query(steps, message, generator))
            records, last = attempt(ask_query, generator, "previous generated cypher query", neo4j.exceptions.CypherSyntaxError,
                                    "Cypher Syntax Error",
                                    "Based on the exception details above, please generate a corrected version of the Cypher query.")
            if last == MAX_ATTEMPTS - 1 or not records:
                print("#" * 100)
                print(f"manually generate cypher query for the following extended metapath: \n {steps}")
                records = run_and_filter_query(state, human_generate_cypher_query(steps, message))
            for record in records:
                report, clues = check_statepath(state, analyzer, record)
                print(report)
                results.append((message, [getattr(x, "id", None) for x in record], report, clues))
    print("*" * 100)
    print(f"ran for {time.time() - started} seconds.")
    print("*" * 100)
    print("close connection")
    meta.close()
    state.close()
    return results
