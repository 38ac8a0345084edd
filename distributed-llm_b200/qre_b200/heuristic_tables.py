"""Pattern tables of the heuristic router.

These regular expressions ARE the behaviour of the reference's `HeuristicRouter`
(src/query_router_engine.py:231-308): category order and pattern order decide which rule fires and
which category name lands in the reasoning string, so they are reproduced verbatim as data.  The C++
batch classifier (csrc/qre_host_features.cu) carries the same tables hand-compiled to literal
alternatives; tests/test_host_features_cpu.py keeps the two in step.
"""

COMPLEX_CATEGORIES = (
    ("code_build_debug", (
        r"\b(write|implement|code|program|script|build|refactor|debug|fix)\b",
        r"\b(traceback|exception|error|segfault|timeout|hanging)\b",
        r"\b(api|flask|fastapi|docker|kubernetes|ssh|tunnel|nginx)\b",
        r"\b(system design|architecture|distributed|scalability|load balanc)\b",
    )),
    ("math_cs_theory", (
        r"\b(prove|lemma|theorem|corollary|derivative|integral|gradient)\b",
        r"\b(time complexity|space complexity|big[- ]o)\b",
        r"\b(dynamic programming|dp|graph|dijkstra|bfs|dfs)\b",
        r"(?:\b|^)a\*(?:\s|$|\W)",
    )),
    ("reasoning_comparison", (
        r"\b(compare|contrast|difference between|pros and cons|vs\.?|versus)\b",
        r"\b(evaluate|assess|critique|analyze)\b",
    )),
    ("long_form_generation", (
        r"\b(essay|report|proposal|research paper|literature review|methodology)\b",
        r"\b(comprehensive|in-depth|step[- ]by[- ]step|detailed)\b",
        r"\b(summarize|synthesis)\b.*\b(everything|all|so far|entire)\b",
        r"\b(transcript|debate|dialogue|format as json|markdown table)\b",
    )),
    ("data_engineering", (
        r"\b(etl|pipeline|spark|hadoop|presto|sql|csv|excel|dataframe|dataset)\b",
        r"\b(deduplicate|normalize|clean|transform|parse|extract)\b",
    )),
    ("medical_analysis", (
        r"\b(symptom|diagnosis|treatment|therapy|prognosis|chronic|severe)\b",
        r"\b(pain|migraine|dizziness|fatigue|nausea|inflammation|anxiety|depression)\b",
        r"\b(dietary|meal|training|exercise|recovery)\b.*\b(plan|schedule|regimen)\b",
        r"\b(mental health|psycholog|counseling|therapist|physician|hospital)\b",
    )),
    ("context_heavy", (
        r"\b(using (all|the) (context|history|above)|based on (the|our) (conversation|context))\b",
        r"\b(continue|expand|build on|follow up)\b.*\b(previous|earlier|above)\b",
    )),
)

SIMPLE_CATEGORIES = (
    ("greeting", (
        r"\b(hi|hello|hey|yo|sup)\b",
        r"\bgood (morning|afternoon|evening)\b",
        r"\b(thanks|thank you)\b",
    )),
    ("general_knowledge", (
        r"\b(what is|who is|where is|when is|when did|how many|capital of)\b",
        r"\b(tell me a joke|fun fact|random fact)\b",
        r"\b(how to|how do i|can you tell me)\b",
    )),
    ("wellness_tips", (
        r"\b(benefits? of|tips? for|advice on)\b",
        r"\b(daily intake|how often|how much)\b",
        r"\b(healthy|good)\b.*\b(habit|routine|lifestyle)\b",
    )),
    ("short_definition", (
        r"\b(define|meaning of|definition of)\b",
        r"\bwhat does\b.*\bmean\b",
    )),
    ("tiny_math", (
        r"^\s*\d+\s*[\+\-\*/]\s*\d+\s*\??\s*$",
        r"^\s*what is\s+\d+\s*[\+\-\*/]\s*\d+\s*\??\s*$",
    )),
)

CODE_MARKERS = (
    "```", "def ", "class ", "import ", "Traceback", "Exception", "ModuleNotFoundError", "SELECT ", "WITH ", "FROM ", "JOIN ", ";", "{", "}", "->", "::", "==", "!=",
)
